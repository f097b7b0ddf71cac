"""Golden vectors for the CBF/CLF constraint block and the obs/state maps, produced by the REFERENCE's own
``SAC_CBF_CLF.get_policy_loss_2`` and ``DynamicsModel`` (imported by path from /root/reference in the build
container; the agent object is created with ``object.__new__`` and given exactly the attributes the method
reads, because its constructor needs gym).  Run here:  ``python tests/golden/make_golden_constraints.py``.
"""
import importlib
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
from oracle import torchdiffeq_restated  # noqa: E402
from make_golden import REF, VARIANTS, _Box  # noqa: E402


def import_variant(tag):
    """Import <variant>/sac_cbf_clf/{model,sac_cbf_clf,dynamics}.py as the package ``sac_cbf_clf``."""
    for k in [k for k in sys.modules if k == "sac_cbf_clf" or k.startswith("sac_cbf_clf.")]:
        del sys.modules[k]
    stub = types.ModuleType("torchdiffeq")
    stub.odeint = torchdiffeq_restated.odeint
    sys.modules["torchdiffeq"] = stub
    vdir = os.path.join(REF, VARIANTS[tag])
    sys.path.insert(0, vdir)
    try:
        model = importlib.import_module("sac_cbf_clf.model")
        model.device = torch.device("cpu")
        agent = importlib.import_module("sac_cbf_clf.sac_cbf_clf")
        dyn = importlib.import_module("sac_cbf_clf.dynamics")
    finally:
        sys.path.remove(vdir)
    return model, agent, dyn


class _Env:
    pass


class _Args:
    cuda = False


def _perturb_biases(net, gen):
    for p in net.parameters():
        if p.ndim == 1:
            p.data.normal_(0, 0.1, generator=gen)


def _with_eps(eps, fn):
    import torch.distributions.normal as tdn
    orig = tdn._standard_normal
    tdn._standard_normal = lambda shape, dtype, device: eps.to(dtype)
    try:
        return fn()
    finally:
        tdn._standard_normal = orig


def _agent(agent_mod, env, **attrs):
    a = object.__new__(agent_mod.SAC_CBF_CLF)
    a.env, a.device, a.solver = env, torch.device("cpu"), "euler"
    a.cost_limit, a.augmented_term, a.augmented_ratio = 0.0, 1.0, 1.0005
    a.Lagrangian_multiplier_update_interval = 8
    for k, v in attrs.items():
        setattr(a, k, v)
    a.lambda_values = [0.0] * a.num_constraints
    return a


def _sd(m):
    return {k: v.clone() for k, v in m.state_dict().items()}


def _run_twice(action, call):
    losses, grads = [], []
    for updates in (8, 9):
        if action.grad is not None:
            action.grad = None
        loss = call(updates)
        loss.backward()
        losses.append(loss.detach().clone())
        grads.append(action.grad.clone())
    return losses, grads


def case_U(B=128, backup=False):
    model, agent_mod, dyn_mod = import_variant("U")
    gen = torch.Generator().manual_seed(321)
    torch.manual_seed(3)
    env = _Env()
    env.dynamics_mode, env.dt, env.hazards_radius = "Unicycle", 0.02, 0.5
    env.hazards_locations = np.array([[0., 0.], [0., 1.], [0., -1.], [-1., 1.], [-1., -1.], [1., -1.], [1., 1.]]) * 1.5
    node = model.NeuralODEModel(3, 3, 6)
    lya = model.LyaNetwork(2, 256)
    _perturb_biases(lya, gen)
    a = _agent(agent_mod, env, gamma_b=50.0, batch_size=B, num_cbfs=7, num_constraints=8, l_p=0.03,
               neural_ode_model=node, lyapunovNet=lya)
    r = lambda *s: torch.rand(*s, generator=gen)  # noqa: E731
    # positions near the hazards so that several CBF rows are active
    state = torch.stack([r(B) * 4 - 2, r(B) * 4 - 2, (r(B) * 2 - 1) * np.pi], 1)
    action = torch.stack([r(B) * 7 - 3.5, r(B) * 24 - 12], 1).requires_grad_(True)
    obs = torch.stack([state[:, 0], state[:, 1], torch.cos(state[:, 2]), torch.sin(state[:, 2]), r(B), r(B), r(B)], 1)
    dm = dyn_mod.DynamicsModel(env, _Args())
    state_from_obs = dm.get_state(obs)
    V = lya(agent_mod.l_p * torch.stack([torch.cos(state[:, 2]), torch.sin(state[:, 2])], 1) + state[:, :2]).detach()
    if backup:   # U sac_cbf_clf.py:532-640: CBF rows only, separate multipliers
        a.backup_lambda_values = [0.0] * a.num_cbfs
        losses, grads = _run_twice(action, lambda u: a.backup_get_policy_loss_2(state, action, u))
        return {"tag": "U", "B": B, "node": _sd(node), "state": state, "action": action.detach(), "losses": losses,
                "grad_action": grads, "lambda_after": [float(v) for v in a.backup_lambda_values],
                "augmented_term_after": a.augmented_term,
                "consts": {"dt": 0.02, "gamma_b": 50.0, "hazards": env.hazards_locations, "hazards_radius": 0.5,
                           "l_p": 0.03}}
    losses, grads = [], []
    for updates in (8, 9):
        if action.grad is not None:
            action.grad = None
        loss = a.get_policy_loss_2(state, action, V, updates)
        loss.backward()
        losses.append(loss.detach().clone())
        grads.append(action.grad.clone())
    return {"tag": "U", "B": B, "node": _sd(node), "lya": _sd(lya), "state": state, "action": action.detach(), "V": V,
            "obs": obs, "state_from_obs": state_from_obs, "losses": losses, "grad_action": grads,
            "lambda_after": [float(v) for v in a.lambda_values], "augmented_term_after": a.augmented_term,
            "consts": {"dt": 0.02, "gamma_b": 50.0, "hazards": env.hazards_locations, "hazards_radius": 0.5,
                       "l_p": 0.03}}


def case_C(B=256, backup=False):
    model, agent_mod, dyn_mod = import_variant("C")
    gen = torch.Generator().manual_seed(322)
    torch.manual_seed(4)
    env = _Env()
    env.dynamics_mode, env.dt = "SimulatedCars", 0.02
    node = model.NeuralODEModel(12, 10)
    lya = model.LyaNetwork(4, 256)
    pol = model.GaussianPolicy(10, 1, 256, _Box([-3.0], [3.0]))
    _perturb_biases(lya, gen)
    _perturb_biases(pol, gen)
    a = _agent(agent_mod, env, gamma_b=0.5, batch_size=B, num_cbfs=2, num_constraints=3, neural_ode_model=node,
               lyapunovNet=lya, policy=pol)
    r = lambda *s: torch.rand(*s, generator=gen)  # noqa: E731
    # gaps near 4.5 so that the CBF rows change sign over the batch
    pos = torch.tensor([42.0, 36, 31, 26, 21]) + (r(B, 5) * 4 - 2)
    vel = 3 + 0.5 * torch.randn(B, 5, generator=gen)
    state = torch.stack([pos, vel], 2).reshape(B, 10)
    action = (r(B, 1) * 6 - 3).requires_grad_(True)
    tk, tk1 = r(B, 1) * 6, r(B, 1) * 6
    eps = torch.randn(B, 1, generator=gen)
    dm = dyn_mod.DynamicsModel(env, _Args())
    obs = dm.get_obs(state.unsqueeze(-1)).squeeze(-1)
    state_from_obs = dm.get_state(obs)
    # V(x_t) is an argument of the method: jitter it so the CLF row changes sign over the batch (with an
    # all-inactive CLF row the reference's own ratio is inf and its loss NaN)
    lv = lya(state[:, 4:8]).detach()
    V = lv + 0.3 * lv.abs().mean() * torch.randn(B, 1, generator=gen)
    if backup:   # C sac_cbf_clf.py:557-681: CBF rows only, the BACKUP policy re-samples u_{t+1}
        bpol = model.GaussianPolicy(10, 1, 256, _Box([-3.0], [3.0]))
        _perturb_biases(bpol, gen)
        a.backup_policy, a.backup_lambda_values = bpol, [0.0] * a.num_cbfs
        losses, grads = _run_twice(action, lambda u: _with_eps(
            eps, lambda: a.backup_get_policy_loss_2(dm, state, action, tk, tk1, u)))
        return {"tag": "C", "B": B, "node": _sd(node), "policy": _sd(bpol), "state": state, "action": action.detach(),
                "t": tk, "t_next": tk1, "eps": eps, "losses": losses, "grad_action": grads,
                "lambda_after": [float(v) for v in a.backup_lambda_values], "augmented_term_after": a.augmented_term,
                "consts": {"dt": 0.02, "gamma_b": 0.5}}
    losses, grads = [], []
    for updates in (8, 9):
        if action.grad is not None:
            action.grad = None
        loss = _with_eps(eps, lambda: a.get_policy_loss_2(dm, state, action, V, tk, tk1, updates))
        loss.backward()
        losses.append(loss.detach().clone())
        grads.append(action.grad.clone())
    return {"tag": "C", "B": B, "node": _sd(node), "lya": _sd(lya), "policy": _sd(pol), "state": state,
            "action": action.detach(), "V": V, "t": tk, "t_next": tk1, "eps": eps, "obs": obs,
            "state_from_obs": state_from_obs, "losses": losses, "grad_action": grads,
            "lambda_after": [float(v) for v in a.lambda_values], "augmented_term_after": a.augmented_term,
            "consts": {"dt": 0.02, "gamma_b": 0.5}}


def case_BU(B=128):
    model, agent_mod, dyn_mod = import_variant("BU")
    gen = torch.Generator().manual_seed(323)
    torch.manual_seed(5)
    env = _Env()
    env.dynamics_mode, env.dt = "Unicycle", 0.02
    node = model.NeuralODEModel(3, 3, 6)
    lya = model.LyaNetwork(2, 256)
    bar = model.BarrierNetwork(7, 2, 256)
    pol = model.GaussianPolicy(7, 2, 256, _Box([-3.5, -12.0], [3.5, 12.0]))
    for n in (lya, bar, pol):
        _perturb_biases(n, gen)
    a = _agent(agent_mod, env, gamma_b=5.0, batch_size=B, num_cbfs=1, num_constraints=2, l_p=0.03,
               neural_ode_model=node, lyapunovNet=lya, BarrierNet=bar, policy=pol)
    r = lambda *s: torch.rand(*s, generator=gen)  # noqa: E731
    state = torch.stack([r(B) * 6 - 3, r(B) * 6 - 3, (r(B) * 2 - 1) * np.pi], 1)
    action = torch.stack([r(B) * 7 - 3.5, r(B) * 24 - 12], 1).requires_grad_(True)
    eps = torch.randn(B, 2, generator=gen)
    dm = dyn_mod.DynamicsModel(env, _Args())
    obs = dm.get_obs(state.unsqueeze(-1), torch.device("cpu")).squeeze(-1).detach()
    state_from_obs = dm.get_state(obs)
    # get_obs gradient fixture
    sg = state.clone().requires_grad_(True)
    cot = torch.randn(B, 7, generator=gen)
    (dm.get_obs(sg.unsqueeze(-1), torch.device("cpu")).squeeze(-1) * cot).sum().backward()
    V = lya(agent_mod.l_p * torch.stack([torch.cos(state[:, 2]), torch.sin(state[:, 2])], 1) + state[:, :2]).detach()
    losses, grads = [], []
    for updates in (8, 9):
        if action.grad is not None:
            action.grad = None
        loss = _with_eps(eps, lambda: a.get_policy_loss_2(obs, dm, state, action, V, updates))
        loss.backward()
        losses.append(loss.detach().clone())
        grads.append(action.grad.clone())
    return {"tag": "BU", "B": B, "node": _sd(node), "lya": _sd(lya), "barrier": _sd(bar), "policy": _sd(pol),
            "state": state, "action": action.detach(), "V": V, "eps": eps, "obs": obs,
            "state_from_obs": state_from_obs, "obs_cot": cot, "grad_state_get_obs": sg.grad.clone(),
            "losses": losses, "grad_action": grads, "lambda_after": [float(v) for v in a.lambda_values],
            "augmented_term_after": a.augmented_term, "consts": {"dt": 0.02, "gamma_b": 5.0, "l_p": 0.03}}


def case_P_maps(B=64):
    """PVTOL get_state / get_obs only (the P constraint block is a 'next' row)."""
    _, _, dyn_mod = import_variant("P")
    gen = torch.Generator().manual_seed(324)
    env = _Env()
    env.dynamics_mode = "Pvtol"
    dm = dyn_mod.DynamicsModel(env, _Args())
    r = lambda *s: torch.rand(*s, generator=gen)  # noqa: E731
    state7 = torch.stack([r(B) * 14 - 7, r(B) * 12 - 6, (r(B) * 2 - 1) * np.pi, r(B) * 4 - 2, r(B) * 4 - 2,
                          r(B) * 2, r(B) * 14 - 7], 1)
    sg = state7.clone().requires_grad_(True)
    obs = dm.get_obs(sg.unsqueeze(-1), torch.device("cpu")).squeeze(-1)
    cot = torch.randn(B, 11, generator=gen)
    (obs * cot).sum().backward()
    st, st_dyn = dm.get_state(obs.detach())
    return {"tag": "P", "state7": state7, "obs": obs.detach(), "obs_cot": cot, "grad_state_get_obs": sg.grad.clone(),
            "state_from_obs": st, "state_dyn_from_obs": st_dyn}


def case_P(B=256, backup=False):
    """PVTOL: 3-step NODE rollout with two policy re-samples, 5 hazard + 4 box/operator third-order CBF rows + CLF
    (P sac_cbf_clf.py:424-738), run by the reference's own method."""
    model, agent_mod, dyn_mod = import_variant("P")
    gen = torch.Generator().manual_seed(325)
    torch.manual_seed(6)
    env = _Env()
    env.dynamics_mode, env.dt, env.hazards_radius = "Pvtol", 0.02, 0.25
    env.hazard_locations = np.array([[-2.5, -2.5], [-2.5, 2.5], [0.0, -3.5], [0.0, 3.5], [-4.5, 0.0]])
    # the env's own operator_dist = 1.0 and y bounds = +-100 never activate on this state range (the operator
    # closes 70 % of its gap every step): tighter values exercise the sign changes of those four rows
    env.safety_operator_follow, env.operator_dist = 0.7, 0.001
    env.y_max, env.y_min = 13.0, -13.0
    node = model.NeuralODEModel(6, 6, 12)
    lya = model.LyaNetwork(11, 256)
    pol = model.GaussianPolicy(11, 2, 256, _Box([-3.5, -15.0], [3.5, 15.0]))
    for n in (lya, pol):
        _perturb_biases(n, gen)
    a = _agent(agent_mod, env, gamma_b=0.6, batch_size=B, num_cbfs=9, num_constraints=10, neural_ode_model=node,
               lyapunovNet=lya, policy=pol)
    r = lambda *s: torch.rand(*s, generator=gen)  # noqa: E731
    x, y = r(B) * 14 - 7, r(B) * 12 - 6
    # a third of the batch sits inside / at the edge of a hazard so that the obstacle rows change sign
    hz = torch.tensor(env.hazard_locations, dtype=torch.float32)
    k = torch.randint(0, 5, (B,), generator=gen)
    near = r(B) < 0.34
    x = torch.where(near, hz[k, 0] + (r(B) * 0.7 - 0.35), x)
    y = torch.where(near, hz[k, 1] + (r(B) * 0.7 - 0.35), y)
    state = torch.stack([x, y, (r(B) * 2 - 1) * np.pi, r(B) * 4 - 2, r(B) * 4 - 2, r(B) * 2,
                         x + (r(B) * 3 - 1.5)], 1)
    action = torch.stack([r(B) * 7 - 3.5, r(B) * 30 - 15], 1).requires_grad_(True)
    eps = torch.randn(B, 2, generator=gen)
    dm = dyn_mod.DynamicsModel(env, _Args())
    obs = dm.get_obs(state.unsqueeze(-1), torch.device("cpu")).squeeze(-1).detach()
    lv = lya(obs).detach()
    V = lv + 0.3 * lv.abs().mean() * torch.randn(B, 1, generator=gen)
    consts = {"dt": 0.02, "gamma_b": 0.6, "hazards": env.hazard_locations, "hazards_radius": 0.25,
              "safety_operator_follow": 0.7, "operator_dist": 0.001, "y_max": 13.0, "y_min": -13.0}
    if backup:   # P sac_cbf_clf.py:740-1045: CBF rows only, the BACKUP policy re-samples u_{t+1}, u_{t+2}
        bpol = model.GaussianPolicy(11, 2, 256, _Box([-3.5, -15.0], [3.5, 15.0]))
        _perturb_biases(bpol, gen)
        a.backup_policy, a.backup_lambda_values = bpol, [0.0] * a.num_cbfs
        # P keeps a separate rho for the backup loss and updates its multipliers every
        # Lagrangian_multiplier_update_interval * backup_update_interval steps (:1025-1034)
        a.backup_update_interval, a.backup_augmented_term = 1, 1.0
        losses, grads = _run_twice(action, lambda u: _with_eps(
            eps, lambda: a.backup_get_policy_loss_2(dm, state, action, u, state[:, :6])))
        return {"tag": "P", "B": B, "node": _sd(node), "policy": _sd(bpol), "state": state, "action": action.detach(),
                "eps": eps, "losses": losses, "grad_action": grads,
                "lambda_after": [float(v) for v in a.backup_lambda_values],
                "augmented_term_after": a.backup_augmented_term, "consts": consts}
    losses, grads = [], []
    for updates in (8, 9):
        if action.grad is not None:
            action.grad = None
        loss = _with_eps(eps, lambda: a.get_policy_loss_2(dm, state, action, V, updates, state[:, :6]))
        loss.backward()
        losses.append(loss.detach().clone())
        grads.append(action.grad.clone())
    return {"tag": "P", "B": B, "node": _sd(node), "lya": _sd(lya), "policy": _sd(pol), "state": state,
            "action": action.detach(), "V": V, "eps": eps, "obs": obs, "losses": losses, "grad_action": grads,
            "lambda_after": [float(v) for v in a.lambda_values], "augmented_term_after": a.augmented_term,
            "consts": {"dt": 0.02, "gamma_b": 0.6, "hazards": env.hazard_locations, "hazards_radius": 0.25,
                       "safety_operator_follow": 0.7, "operator_dist": 0.001, "y_max": 13.0, "y_min": -13.0}}


def case_BP(B=256):
    """Learned-barrier PVTOL (BP sac_cbf_clf.py:360-472): one NODE step + operator follow, barrier net on the
    predicted observation with the re-sampled action, CLF row; ratio floor 0.002."""
    model, agent_mod, dyn_mod = import_variant("BP")
    gen = torch.Generator().manual_seed(326)
    torch.manual_seed(7)
    env = _Env()
    env.dynamics_mode, env.dt, env.safety_operator_follow = "Pvtol", 0.02, 0.7
    node = model.NeuralODEModel(6, 6, 12)
    lya = model.LyaNetwork(11, 256)
    bar = model.BarrierNetwork(11, 2, 256)
    pol = model.GaussianPolicy(11, 2, 256, _Box([-3.5, -15.0], [3.5, 15.0]))
    for n in (lya, bar, pol):
        _perturb_biases(n, gen)
    a = _agent(agent_mod, env, gamma_b=0.6, batch_size=B, num_cbfs=1, num_constraints=2, neural_ode_model=node,
               lyapunovNet=lya, BarrierNet=bar, policy=pol)
    r = lambda *s: torch.rand(*s, generator=gen)  # noqa: E731
    x = r(B) * 14 - 7
    state = torch.stack([x, r(B) * 12 - 6, (r(B) * 2 - 1) * np.pi, r(B) * 4 - 2, r(B) * 4 - 2, r(B) * 2,
                         x + (r(B) * 3 - 1.5)], 1)
    action = torch.stack([r(B) * 7 - 3.5, r(B) * 30 - 15], 1).requires_grad_(True)
    eps = torch.randn(B, 2, generator=gen)
    dm = dyn_mod.DynamicsModel(env, _Args())
    obs = dm.get_obs(state.unsqueeze(-1), torch.device("cpu")).squeeze(-1).detach()
    lv = lya(obs).detach()
    V = lv + 0.3 * lv.abs().mean() * torch.randn(B, 1, generator=gen)
    losses, grads = _run_twice(action, lambda u: _with_eps(
        eps, lambda: a.get_policy_loss_2(obs, dm, state, action, V, u, state[:, :6])))
    return {"tag": "BP", "B": B, "node": _sd(node), "lya": _sd(lya), "barrier": _sd(bar), "policy": _sd(pol),
            "state": state, "action": action.detach(), "V": V, "eps": eps, "obs": obs, "losses": losses,
            "grad_action": grads, "lambda_after": [float(v) for v in a.lambda_values],
            "augmented_term_after": a.augmented_term,
            "consts": {"dt": 0.02, "gamma_b": 0.6, "safety_operator_follow": 0.7}}


def main():
    for name, fn in (("U", case_U), ("C", case_C), ("BU", case_BU), ("Pmaps", case_P_maps), ("P", case_P),
                     ("BP", case_BP)):
        torch.save(fn(), os.path.join(HERE, "constraints_%s.pt" % name))
        print("constraints", name, "ok")
    for name, fn in (("U", case_U), ("C", case_C), ("P", case_P)):
        torch.save(fn(backup=True), os.path.join(HERE, "constraints_backup_%s.pt" % name))
        print("backup constraints", name, "ok")


if __name__ == "__main__":
    main()
