"""Shared builders: rebuild the nets of a constraints_*.pt golden case for either the oracle (CPU restatement)
or the product (CUDA host mirror), and evaluate the two-call loss sequence the golden run recorded."""
import torch


def build(g, lib, device):
    """lib = oracle.nlbac_ref (AffineNODE/BlackBoxNODE) or nlbac_b200 (NeuralODEModel)."""
    tag = g["tag"]
    is_oracle = hasattr(lib, "AffineNODE")
    if tag in ("U", "BU"):
        node = lib.AffineNODE(3, 3, 6) if is_oracle else lib.NeuralODEModel(3, 3, 6)
        lya = lib.LyaNetwork(2, 256)
    elif tag in ("P", "BP"):
        node = lib.AffineNODE(6, 6, 12) if is_oracle else lib.NeuralODEModel(6, 6, 12)
        lya = lib.LyaNetwork(11, 256)
    else:
        node = lib.BlackBoxNODE(12, 10) if is_oracle else lib.NeuralODEModel(12, 10)
        lya = lib.LyaNetwork(4, 256)
    node.load_state_dict(g["node"])
    nets = {"node": node.to(device)}
    if "lya" in g:        # the backup-loss goldens have no Lyapunov net
        lya.load_state_dict(g["lya"])
        nets["lya"] = lya.to(device)
    if "policy" in g:
        n_o, n_u, hi = (10, 1, [3.0]) if tag == "C" else ((11, 2, [3.5, 15.0]) if tag in ("P", "BP") else (7, 2, [3.5, 12.0]))
        if is_oracle:
            pol = lib.GaussianPolicy(n_o, n_u, 256, hi, [-h for h in hi])
        else:
            import numpy as np

            class Box:
                high = np.array(hi, dtype="float32")
                low = -np.array(hi, dtype="float32")
            pol = lib.GaussianPolicy(n_o, n_u, 256, Box)
        pol.load_state_dict(g["policy"])
        nets["policy"] = pol.to(device)
    if "barrier" in g:
        bar = lib.BarrierNetwork(11 if tag == "BP" else 7, 2, 256)
        bar.load_state_dict(g["barrier"])
        nets["barrier"] = bar.to(device)
    return nets


def run_sequence(g, nets, C, device, dyn=None):
    """C = oracle.constraints or nlbac_b200.constraints; returns (losses, grad_action list, state object)."""
    tag, B, k = g["tag"], g["B"], g["consts"]
    is_oracle = not hasattr(C, "ConstraintState") or "batch_size" not in C.ConstraintState.__init__.__code__.co_varnames
    state = g["state"].to(device)
    V = g["V"].to(device)
    action = g["action"].clone().to(device).requires_grad_(True)
    if tag == "U":
        st = C.ConstraintState(8) if is_oracle else C.ConstraintState(8, B)
    elif tag == "C":
        st = (C.ConstraintState(3, lambda_max=300.0, ratio_floor=0.002) if is_oracle
              else C.ConstraintState(3, B, lambda_max=300.0, ratio_floor=0.002))
    elif tag == "P":      # P sac_cbf_clf.py:707-708 (floor 0.002), :718 (clamp 400)
        st = (C.ConstraintState(10, ratio_floor=0.002) if is_oracle else C.ConstraintState(10, B, ratio_floor=0.002))
    elif tag == "BP":     # BP sac_cbf_clf.py:441-447 (ratio with the 0.002 floor)
        st = C.ConstraintState(2, ratio_floor=0.002) if is_oracle else C.ConstraintState(2, B, ratio_floor=0.002)
    else:
        st = C.ConstraintState(2, use_ratio=False) if is_oracle else C.ConstraintState(2, B, use_ratio=False)
    eps = g["eps"].to(device) if "eps" in g else None
    losses, grads = [], []
    for updates in (8, 9):
        action.grad = None
        if tag == "BP":
            obs = g["obs"].to(device)
            kw = dict(dt=k["dt"], gamma_b=k["gamma_b"], batch_size=B, eps_next=eps)
            if is_oracle:
                req = C.barrier_pvtol_required_matrix(nets["node"], nets["lya"], nets["barrier"], nets["policy"], obs,
                                                      state, action, V, follow=k["safety_operator_follow"], **kw)
            else:
                req = C.barrier_pvtol_required_matrix(nets["node"], nets["lya"], nets["barrier"], nets["policy"], dyn,
                                                      obs, state, action, V,
                                                      safety_operator_follow=k["safety_operator_follow"], **kw)
        elif tag == "U":
            kw = dict(dt=k["dt"], gamma_b=k["gamma_b"], hazards_radius=k["hazards_radius"], l_p=k["l_p"], batch_size=B)
            if is_oracle:
                req = C.unicycle_required_matrix(nets["node"], nets["lya"], state, action, V, hazards=k["hazards"], **kw)
            else:
                req = C.unicycle_required_matrix(nets["node"], nets["lya"], state, action, V,
                                                 hazards_locations=k["hazards"], **kw)
        elif tag == "C":
            t, t1 = g["t"].to(device), g["t_next"].to(device)
            if is_oracle:
                req = C.cars_required_matrix(nets["node"], nets["lya"], nets["policy"], state, action, V, t, t1,
                                             dt=k["dt"], gamma_b=k["gamma_b"], batch_size=B, eps_next=eps)
            else:
                req = C.cars_required_matrix(nets["node"], nets["lya"], nets["policy"], dyn, state, action, V, t, t1,
                                             dt=k["dt"], gamma_b=k["gamma_b"], batch_size=B, eps_next=eps)
        elif tag == "P":
            kw = dict(dt=k["dt"], gamma_b=k["gamma_b"], hazards_radius=k["hazards_radius"],
                      operator_dist=k["operator_dist"], y_max=k["y_max"], y_min=k["y_min"], batch_size=B,
                      eps_next=eps, eps_next_next=eps)
            if is_oracle:
                req = C.pvtol_required_matrix(nets["node"], nets["lya"], nets["policy"], state, action, V,
                                              hazards=k["hazards"], follow=k["safety_operator_follow"], **kw)
            else:
                req = C.pvtol_required_matrix(nets["node"], nets["lya"], nets["policy"], dyn, state, action, V,
                                              hazard_locations=k["hazards"],
                                              safety_operator_follow=k["safety_operator_follow"], **kw)
        else:
            obs = g["obs"].to(device)
            if is_oracle:
                req = C.barrier_unicycle_required_matrix(nets["node"], nets["lya"], nets["barrier"], nets["policy"], obs,
                                                         state, action, V, dt=k["dt"], gamma_b=k["gamma_b"],
                                                         l_p=k["l_p"], batch_size=B, eps_next=eps)
            else:
                req = C.barrier_unicycle_required_matrix(nets["node"], nets["lya"], nets["barrier"], nets["policy"], dyn,
                                                         obs, state, action, V, dt=k["dt"], gamma_b=k["gamma_b"],
                                                         l_p=k["l_p"], batch_size=B, eps_next=eps)
        loss = C.augmented_lagrangian(req, st, updates)
        loss.backward()
        losses.append(loss.detach().reshape(-1).cpu())
        grads.append(action.grad.detach().cpu().clone())
    return losses, grads, st


def run_backup_sequence(g, nets, C, device, dyn=None):
    """backup_get_policy_loss_2 twice (updates 8, 9) as the constraints_backup_*.pt goldens recorded it: CBF rows
    only, the backup policy re-samples, separate multipliers (P: separate rho)."""
    tag, B, k = g["tag"], g["B"], g["consts"]
    is_oracle = "batch_size" not in C.ConstraintState.__init__.__code__.co_varnames
    state = g["state"].to(device)
    action = g["action"].clone().to(device).requires_grad_(True)
    n_con = {"U": 8, "C": 3, "P": 10}[tag]
    kw_st = {"C": dict(lambda_max=300.0), "P": dict(separate_backup_rho=True, backup_update_interval=1)}.get(tag, {})
    st = C.ConstraintState(n_con, **kw_st) if is_oracle else C.ConstraintState(n_con, B, **kw_st)
    eps = g["eps"].to(device) if "eps" in g else None
    losses, grads = [], []
    for updates in (8, 9):
        action.grad = None
        if tag == "U":
            kw = dict(dt=k["dt"], gamma_b=k["gamma_b"], hazards_radius=k["hazards_radius"], l_p=k["l_p"], batch_size=B)
            if is_oracle:
                req = C.unicycle_required_matrix(nets["node"], None, state, action, None, hazards=k["hazards"], **kw)
            else:
                req = C.unicycle_required_matrix(nets["node"], None, state, action, None,
                                                 hazards_locations=k["hazards"], **kw)
        elif tag == "C":
            t, t1 = g["t"].to(device), g["t_next"].to(device)
            args = (state, action, None, t, t1)
            kw = dict(dt=k["dt"], gamma_b=k["gamma_b"], batch_size=B, eps_next=eps)
            if is_oracle:
                req = C.cars_required_matrix(nets["node"], None, nets["policy"], *args, **kw)
            else:
                req = C.cars_required_matrix(nets["node"], None, nets["policy"], dyn, *args, **kw)
        else:
            kw = dict(dt=k["dt"], gamma_b=k["gamma_b"], hazards_radius=k["hazards_radius"],
                      operator_dist=k["operator_dist"], y_max=k["y_max"], y_min=k["y_min"], batch_size=B,
                      eps_next=eps, eps_next_next=eps)
            if is_oracle:
                req = C.pvtol_required_matrix(nets["node"], None, nets["policy"], state, action, None,
                                              hazards=k["hazards"], follow=k["safety_operator_follow"], **kw)
            else:
                req = C.pvtol_required_matrix(nets["node"], None, nets["policy"], dyn, state, action, None,
                                              hazard_locations=k["hazards"],
                                              safety_operator_follow=k["safety_operator_follow"], **kw)
        loss = C.backup_augmented_lagrangian(req, st, updates)
        loss.backward()
        losses.append(loss.detach().reshape(-1).cpu())
        grads.append(action.grad.detach().cpu().clone())
    return losses, grads, st
