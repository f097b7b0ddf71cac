"""CPU restatement of the reference's CBF/CLF constraint block and obs/state maps.

TEST INFRASTRUCTURE (see oracle/__init__.py) — only tests/, smoke() and bench.py's cpu_baseline may import it.
Plain PyTorch, device-agnostic, differentiable.  Pinned against the reference's own ``get_policy_loss_2`` /
``DynamicsModel`` executed in the build container (tests/golden/make_golden_constraints.py ->
tests/golden/constraints_*.pt -> tests/test_oracle_constraints.py).

Path tags as in nlbac_ref.py (U, C, P, BU, BP).
"""
import math

import numpy as np
import torch

from .torchdiffeq_restated import odeint


# ------------------------------------------------------------------------------------------------------
# DynamicsModel.get_state / get_obs
# ------------------------------------------------------------------------------------------------------
def get_state(env, obs):
    """U dynamics.py:27-69, C :25-67, P :25-73.  Tensor in -> float64 numpy arithmetic -> tensor of the input
    dtype (the reference's ``to_numpy`` / ``to_tensor`` round trip, U utils.py:60-72); autograd history dropped."""
    o = obs.detach().cpu().double().numpy()
    expand = o.ndim == 1
    if expand:
        o = o[None]
    if env == "Unicycle":
        st = np.stack([o[:, 0], o[:, 1], np.arctan2(o[:, 3], o[:, 2])], 1)
    elif env == "SimulatedCars":
        st = o.copy()
        st[:, ::2] *= 100.0
        st[:, 1::2] *= 30.0
    elif env == "Pvtol":
        st = np.stack([o[:, 0], o[:, 1], np.arctan2(o[:, 3], o[:, 2]), o[:, 4], o[:, 5], o[:, 6], o[:, 7]], 1)
    else:
        raise Exception("Unknown dynamics")
    t = torch.from_numpy(st[0] if expand else st).to(obs.dtype).to(obs.device)
    if env == "Pvtol":
        return t, t[..., :6]
    return t


def get_obs(env, state):
    """Differentiable state -> obs on ``[B, n_s]`` (the reference works on ``[B, n_s, 1]``): C dynamics.py:69-96
    (divide by 100 / 30), P :75-153 and BU :73-198 (``[x, y, cos, sin, pass-through..., compass(2), exp(-dist)]``,
    compass = R(theta) (goal - p) / (|.| + 1e-3), goal (4.5, 4.5) for Pvtol and (2.5, 2.5) for Unicycle)."""
    if env == "SimulatedCars":
        scale = torch.tensor([100.0, 30.0] * 5, dtype=state.dtype, device=state.device)
        return state / scale
    goal = 4.5 if env == "Pvtol" else 2.5
    th = state[:, 2]
    c, s = torch.cos(th), torch.sin(th)
    rel = goal - state[:, :2]
    dist = torch.norm(rel, dim=1)
    vx = c * rel[:, 0] + s * rel[:, 1]
    vy = -s * rel[:, 0] + c * rel[:, 1]
    div = torch.sqrt(vx * vx + vy * vy) + 0.001
    cols = [state[:, 0], state[:, 1], c, s] + [state[:, k] for k in range(3, state.shape[1])]
    cols += [vx / div, vy / div, torch.exp(-dist)]
    return torch.stack(cols, 1)


# ------------------------------------------------------------------------------------------------------
# constraint rows -> required_matrix
# ------------------------------------------------------------------------------------------------------
def required_matrix_from_rows(matr, batch_size):
    """U sac_cbf_clf.py:493-504: ``where(matr > 0, matr, 0)`` summed over the batch, divided by ``args.batch_size``."""
    filtered = torch.where(matr > 0, matr, torch.zeros_like(matr))
    return (filtered.sum(0) / batch_size).unsqueeze(-1)


def lookahead(state, l_p):
    th = state[:, 2]
    return torch.stack([state[:, 0] + l_p * torch.cos(th), state[:, 1] + l_p * torch.sin(th)], 1)


def unicycle_required_matrix(node, lya, state, action, V, *, dt, gamma_b, hazards, hazards_radius, l_p, batch_size,
                             solver="euler", gamma_l=1.0):
    """U sac_cbf_clf.py:420-504."""
    rc = 1.05 * hazards_radius
    hz = torch.as_tensor(hazards, dtype=state.dtype, device=state.device)
    ps = lookahead(state, l_p)
    y0 = torch.cat((state, action), -1)
    nxt = odeint(node, y0, torch.tensor([0.0, dt]), method=solver, atol=1e-7, rtol=1e-5)[-1]
    ps_next = lookahead(nxt[:, :3], l_p)
    hs = 0.5 * (((ps[:, None, :] - hz[None]) ** 2).sum(2) - rc ** 2)
    hs_next = 0.5 * (((ps_next[:, None, :] - hz[None]) ** 2).sum(2) - rc ** 2)
    cbf = -((hs_next - hs) / dt) - gamma_b * hs
    if lya is None:   # backup_get_policy_loss_2 (U :532-608): CBF rows only
        return required_matrix_from_rows(cbf, batch_size)
    V_next = lya(ps_next)
    lya_term = (V_next - V) / dt + gamma_l * V
    return required_matrix_from_rows(torch.cat((cbf, lya_term), 1), batch_size)


def cars_required_matrix(node, lya, policy, state, action, V, t_batch, next_t_batch, *, dt, gamma_b, batch_size,
                         solver="euler", gamma_l=0.15, eps_next=None):
    """C sac_cbf_clf.py:424-523."""
    r = 4.5
    tspan = torch.tensor([0.0, dt])
    x1 = odeint(node, torch.cat((state, action, t_batch), -1), tspan, method=solver, atol=1e-7, rtol=1e-5)[-1][:, :10]
    obs1 = get_obs("SimulatedCars", x1).detach()
    pi_next = policy.sample(obs1, eps_next)[0].detach()
    x2 = odeint(node, torch.cat((x1, pi_next, next_t_batch), -1), tspan, method=solver, atol=1e-7, rtol=1e-5)[-1][:, :10]
    rows = []
    for ia, ib in ((4, 6), (6, 8)):
        h0, h1, h2 = (x[:, ia:ia + 1] - x[:, ib:ib + 1] - r for x in (state, x1, x2))
        l1 = h1 - h0 + gamma_b * h0
        l2 = h2 - h1 + gamma_b * h1
        rows.append(-(l2 - l1) - gamma_b * l1)
    if lya is None:   # backup_get_policy_loss_2 (C :557-645): CBF rows only, `policy` is the backup policy
        return required_matrix_from_rows(torch.cat(rows, 1), batch_size)
    V_next = lya(x1[:, 4:8])
    lya_term = (V_next - V) + gamma_l * V
    return required_matrix_from_rows(torch.cat(rows + [lya_term], 1), batch_size)


def barrier_unicycle_required_matrix(node, lya, barrier, policy, obs, state, action, V, *, dt, gamma_b, l_p,
                                     batch_size, solver="euler", gamma_l=1.0, eps_next=None):
    """BU sac_cbf_clf.py:368-446."""
    nxt = odeint(node, torch.cat((state, action), -1), torch.tensor([0.0, dt]), method=solver, atol=1e-7, rtol=1e-5)[-1]
    pred = nxt[:, :3]
    V_next = lya(lookahead(pred, l_p))
    lya_term = (V_next - V) / dt + gamma_l * V
    b_now = barrier(obs, action).detach()
    next_obs = get_obs("Unicycle", pred)
    pi_next = policy.sample(next_obs.detach(), eps_next)[0].detach()
    b_next = barrier(next_obs, pi_next)
    barrier_term = -(b_next - b_now) - gamma_b * b_now
    return required_matrix_from_rows(torch.cat((barrier_term, lya_term), 1), batch_size)


def barrier_pvtol_required_matrix(node, lya, barrier, policy, obs, state, action, V, *, dt, gamma_b, follow,
                                  batch_size, solver="euler", gamma_l=0.1, eps_next=None):
    """BP sac_cbf_clf.py:373-438."""
    nxt = odeint(node, torch.cat((state[:, :6], action), -1), torch.tensor([0.0, dt]), method=solver, atol=1e-7,
                 rtol=1e-5)[-1][:, :6]
    op = state[:, 6] + follow * (nxt[:, 0] - state[:, 6])
    next_obs = get_obs("Pvtol", torch.cat((nxt, op[:, None]), 1))
    lya_term = (lya(next_obs) - V) / 1.0 + gamma_l * V
    pi_next = policy.sample(next_obs.detach(), eps_next)[0].detach()
    b_now = barrier(obs, action).detach()
    b_next = barrier(next_obs, pi_next)
    barrier_term = -(b_next - b_now) - gamma_b * b_now
    return required_matrix_from_rows(torch.cat((barrier_term, lya_term), 1), batch_size)


def third_order_cbf(h0, h1, h2, h3, gamma_b):
    """P sac_cbf_clf.py:581-587 (same expression for every P row): a discrete third-order CBF condition."""
    t1 = h3 - h2 + gamma_b * h2
    t2 = h2 - h1 + gamma_b * h1
    t3 = h1 - h0 + gamma_b * h0
    return -(t1 - t2 + gamma_b * t2 - (t2 - t3 + gamma_b * t3) + gamma_b * (t2 - t3 + gamma_b * t3))


def pvtol_rollout(node, policy, state, action, *, dt, follow, solver="euler", eps_next=None, eps_next_next=None):
    """P sac_cbf_clf.py:452-552: three chained NODE predictions of the 6 dynamic states; the 7th state (the safety
    operator's x) follows ``op + follow * (x_next - op)``; the policy is re-sampled on the detached predicted
    observation between steps.  Returns the three 7-column states."""
    tspan = torch.tensor([0.0, dt])

    def step(s7_prev, dyn_prev, u):
        nxt = odeint(node, torch.cat((dyn_prev, u), -1), tspan, method=solver, atol=1e-7, rtol=1e-5)[-1][:, :6]
        op = s7_prev[:, 6] + follow * (nxt[:, 0] - s7_prev[:, 6])
        return torch.cat((nxt, op[:, None]), 1), nxt

    s1, d1 = step(state, state[:, :6], action)
    u1 = policy.sample(get_obs("Pvtol", s1).detach(), eps_next)[0].detach()
    s2, d2 = step(s1, d1, u1)
    u2 = policy.sample(get_obs("Pvtol", s2).detach(), eps_next_next)[0].detach()
    s3, _ = step(s2, d2, u2)
    return s1, s2, s3


def pvtol_required_matrix(node, lya, policy, state, action, V, *, dt, gamma_b, hazards, hazards_radius, follow,
                          operator_dist, y_max, y_min, batch_size, solver="euler", gamma_l=0.1, eps_next=None,
                          eps_next_next=None):
    """P sac_cbf_clf.py:424-697: rows = 5 hazards, operator (two sides), y_max, y_min, CLF."""
    rc = 1.2 * hazards_radius
    od = 0.9 * operator_dist
    delta_y = 10.0
    hz = torch.as_tensor(hazards, dtype=state.dtype, device=state.device)
    s1, s2, s3 = pvtol_rollout(node, policy, state, action, dt=dt, follow=follow, solver=solver, eps_next=eps_next,
                               eps_next_next=eps_next_next)
    seq = (state, s1, s2, s3)
    hs = [0.5 * (((s[:, None, :2] - hz[None]) ** 2).sum(2) - rc ** 2) for s in seq]
    rows = [third_order_cbf(*hs, gamma_b)]
    for f in (lambda s: s[:, 0:1] - s[:, 6:7] + od, lambda s: -s[:, 0:1] + s[:, 6:7] + od,
              lambda s: -s[:, 1:2] + y_max - delta_y, lambda s: s[:, 1:2] - y_min - delta_y):
        rows.append(third_order_cbf(*[f(s) for s in seq], gamma_b))
    if lya is None:   # backup_get_policy_loss_2 (P :740-1010): CBF rows only, `policy` is the backup policy
        return required_matrix_from_rows(torch.cat(rows, 1), batch_size)
    V_next = lya(get_obs("Pvtol", s1))
    lya_term = (V_next - V) / 1.0 + gamma_l * V
    return required_matrix_from_rows(torch.cat(rows + [lya_term], 1), batch_size)


class ConstraintState:
    """lambda / rho bookkeeping of the agent (U sac_cbf_clf.py:55-62)."""

    def __init__(self, num_constraints, lambda_max=400.0, ratio_floor=None, use_ratio=True, interval=8,
                 backup_update_interval=1, separate_backup_rho=False):
        self.lambda_values = [0.0] * num_constraints
        self.backup_lambda_values = [0.0] * (num_constraints - 1)
        self.num_constraints = num_constraints
        self.augmented_term, self.augmented_ratio, self.cost_limit = 1.0, 1.0005, 0.0
        self.backup_augmented_term = 1.0
        self.lambda_max, self.ratio_floor, self.use_ratio, self.interval = lambda_max, ratio_floor, use_ratio, interval
        self.backup_update_interval, self.separate_backup_rho = backup_update_interval, separate_backup_rho


def backup_augmented_lagrangian(req, st, updates):
    """U sac_cbf_clf.py:611-640, C :657-681 (shared rho); P :1022-1045 (its own rho, multipliers updated every
    ``interval * backup_update_interval`` steps).  No ratio: every row enters with its own multiplier."""
    rho_name = "backup_augmented_term" if st.separate_backup_rho else "augmented_term"
    rc = req.detach()
    if updates % (st.interval * st.backup_update_interval) == 0:
        for i in range(len(st.backup_lambda_values)):
            st.backup_lambda_values[i] = torch.clamp(st.backup_lambda_values[i] + getattr(st, rho_name) * rc[i], 0.01,
                                                     st.lambda_max)
    setattr(st, rho_name, min(getattr(st, rho_name) * st.augmented_ratio, 200))
    rho = getattr(st, rho_name)
    loss = 0.0
    for i in range(req.shape[0]):
        c = req[i] - st.cost_limit
        loss = loss + float(st.backup_lambda_values[i]) * c + rho / 2.0 * c * c
    return loss


def augmented_lagrangian(req, st, updates):
    """U sac_cbf_clf.py:506-530; C :525-555 (ratio floor 0.002, clamp 300); BU :462-477 (no ratio)."""
    ratio = 1.0
    if st.use_ratio:
        other_mean = torch.abs(torch.mean(req[:-1, :] - st.cost_limit))
        lya_c = torch.abs(req[-1, :] - st.cost_limit)
        ratio = float(other_mean / lya_c)
        if st.ratio_floor is not None and ratio < st.ratio_floor:
            ratio = st.ratio_floor
    rc = req.detach()
    if updates % st.interval == 0:
        for i in range(st.num_constraints):
            st.lambda_values[i] = torch.clamp(st.lambda_values[i] + st.augmented_term * rc[i], 0.01, st.lambda_max)
    st.augmented_term = min(st.augmented_term * st.augmented_ratio, 200)
    rho = st.augmented_term
    loss = 0.0
    for i in range(req.shape[0] - 1):
        c = req[i] - st.cost_limit
        loss = loss + float(st.lambda_values[i]) * c + rho / 2.0 * c * c
    c = req[-1] - st.cost_limit
    return loss + float(st.lambda_values[-1]) * ratio * c + ratio * ratio * rho / 2.0 * c * c


def chained_rollout(node, policy, state, action, horizon, *, dt, eps=None, solver="euler", obs_fn=None):
    """Restatement of the chained-prediction pattern of C sac_cbf_clf.py:437-459 / P :459-552 at horizon H (config 4):
    predict, re-sample the policy on the detached predicted observation, predict again."""
    n_s = state.shape[1]
    x, u, out = state, action, []
    for k in range(horizon):
        x = odeint(node, torch.cat((x, u), -1), torch.tensor([0.0, dt]), method=solver, atol=1e-7, rtol=1e-5)[-1][:, :n_s]
        out.append(x)
        if k + 1 < horizon:
            obs = x.detach() if obs_fn is None else obs_fn(x.detach())
            u = policy.sample(obs, None if eps is None else eps[k])[0].detach()
    return torch.stack(out)
