"""``get_policy_loss_2`` — the CBF/CLF augmented-Lagrangian part of the actor loss — on the fused kernels.

Mirrors the bodies of ``SAC_CBF_CLF.get_policy_loss_2`` for Unicycle (U sac_cbf_clf.py:408-530), SimulatedCars
(C :412-555), PVTOL (P :424-738) and the learned-barrier variants (BU :359-477, BP :360-472): one fused NODE prediction per rollout
step (``odeint``), the fused Lyapunov / barrier / policy MLPs, and ONE fused kernel for the constraint rows,
the ``where(>0)`` filter and the batch reduction.  The multiplier / rho bookkeeping is the reference's, kept on
the host exactly as written there (python floats, App. D.3).
"""
import torch

from . import _lib
from .odeint import odeint
from .tail import constraint_sums, lookahead, make_constraint_desc


class ConstraintState:
    """The agent attributes the reference's method reads and mutates (U sac_cbf_clf.py:55-62)."""

    def __init__(self, num_constraints, batch_size, *, lambda_max=400.0, cost_limit=0.0, augmented_term=1.0,
                 augmented_ratio=1.0005, update_interval=8, ratio_floor=None, use_ratio=True,
                 backup_update_interval=1, separate_backup_rho=False):
        self.num_constraints = num_constraints
        self.batch_size = batch_size
        self.lambda_values = [0.0] * num_constraints
        self.backup_lambda_values = [0.0] * (num_constraints - 1)      # backup policy: CBF rows only
        self.backup_augmented_term = augmented_term                    # P keeps its own rho (P :1032-1034)
        self.backup_update_interval = backup_update_interval           # P: args.backup_update_interval (:1025)
        self.separate_backup_rho = separate_backup_rho
        self.lambda_max = lambda_max
        self.cost_limit = cost_limit
        self.augmented_term = augmented_term
        self.augmented_ratio = augmented_ratio
        self.Lagrangian_multiplier_update_interval = update_interval
        self.ratio_floor = ratio_floor
        self.use_ratio = use_ratio


def augmented_lagrangian(required_matrix, st, updates):
    """U sac_cbf_clf.py:506-530 (C :525-555 adds the 0.002 floor; BU :462-477 has no ratio)."""
    ratio = 1.0
    if st.use_ratio:
        other = required_matrix[:-1, :] - st.cost_limit
        other_mean = torch.abs(torch.mean(other))
        lya_component = torch.abs(required_matrix[-1, :] - st.cost_limit)
        ratio = float(other_mean / lya_component)
        if st.ratio_floor is not None and ratio < st.ratio_floor:
            ratio = st.ratio_floor
    req_copy = required_matrix.detach()
    if updates % st.Lagrangian_multiplier_update_interval == 0:
        for i in range(st.num_constraints):
            new_lambda = st.lambda_values[i] + st.augmented_term * req_copy[i]
            st.lambda_values[i] = torch.clamp(new_lambda, 0.01, st.lambda_max)
    st.augmented_term = min(st.augmented_term * st.augmented_ratio, 200)
    rho, cl = st.augmented_term, st.cost_limit
    loss = 0.0
    m = required_matrix.shape[0]
    for i in range(m - 1):
        c = required_matrix[i] - cl
        loss = loss + float(st.lambda_values[i]) * c + rho / 2.0 * c * c
    c = required_matrix[-1] - cl
    loss = loss + float(st.lambda_values[-1]) * ratio * c + ratio * ratio * rho / 2.0 * c * c
    return loss


def backup_augmented_lagrangian(backup_required_matrix, st, updates):
    """The tail of ``backup_get_policy_loss_2``: U sac_cbf_clf.py:611-640, C :657-681 (rho shared with the primary
    loss); P :1022-1045 (its own rho; multipliers every ``interval * backup_update_interval`` updates)."""
    rho_name = "backup_augmented_term" if st.separate_backup_rho else "augmented_term"
    req_copy = backup_required_matrix.detach()
    if updates % (st.Lagrangian_multiplier_update_interval * st.backup_update_interval) == 0:
        for i in range(len(st.backup_lambda_values)):
            new_lambda = st.backup_lambda_values[i] + getattr(st, rho_name) * req_copy[i]
            st.backup_lambda_values[i] = torch.clamp(new_lambda, 0.01, st.lambda_max)
    setattr(st, rho_name, min(getattr(st, rho_name) * st.augmented_ratio, 200))
    rho, cl = getattr(st, rho_name), st.cost_limit
    loss = 0.0
    for i in range(backup_required_matrix.shape[0]):
        c = backup_required_matrix[i] - cl
        loss = loss + float(st.backup_lambda_values[i]) * c + rho / 2.0 * c * c
    return loss


class DeviceConstraintState:
    """The same bookkeeping as ``ConstraintState`` with every scalar ON THE DEVICE (float64, mirroring the Python
    floats of the reference), so that ``augmented_lagrangian_device`` has no host read-back (the reference's
    ``float(ratio)`` / ``float(lambda)`` at U sac_cbf_clf.py:510, 525-528) and the whole actor update can be captured
    in a CUDA graph.  ``set_updates(updates)`` is the only per-call host input: it writes the
    "multipliers are due" flag (``updates % interval == 0``) into a device scalar before the replay."""

    def __init__(self, num_constraints, batch_size, device, *, lambda_max=400.0, cost_limit=0.0, augmented_term=1.0,
                 augmented_ratio=1.0005, update_interval=8, ratio_floor=None, use_ratio=True,
                 backup_update_interval=1, separate_backup_rho=False):
        f64 = dict(dtype=torch.float64, device=device)
        self.num_constraints, self.batch_size = num_constraints, batch_size
        self.lambda_values = torch.zeros(num_constraints, **f64)
        self.augmented_term = torch.full((), float(augmented_term), **f64)
        self.due = torch.zeros((), **f64)
        # backup policy: CBF rows only, its own multipliers; U / C share rho with the primary loss (advanced twice per
        # update), P keeps a separate one and a longer multiplier interval (P :1025-1034)
        self.backup_lambda_values = torch.zeros(num_constraints - 1, **f64)
        self.backup_augmented_term = torch.full((), float(augmented_term), **f64)
        self.backup_due = torch.zeros((), **f64)
        self.backup_update_interval, self.separate_backup_rho = backup_update_interval, separate_backup_rho
        self.lambda_max, self.cost_limit, self.augmented_ratio = lambda_max, cost_limit, augmented_ratio
        self.Lagrangian_multiplier_update_interval = update_interval
        self.ratio_floor, self.use_ratio = ratio_floor, use_ratio

    def set_updates(self, updates):
        self._updates_dev = None
        self.due.fill_(1.0 if updates % self.Lagrangian_multiplier_update_interval == 0 else 0.0)
        self.backup_due.fill_(1.0 if updates % (self.Lagrangian_multiplier_update_interval *
                                                 self.backup_update_interval) == 0 else 0.0)

    def set_updates_device(self, updates_dev):
        """Same from a device-resident update counter (int64 tensor): no host input at all, capturable.  Nothing is
        launched here: the fused Lagrangian kernel reads the counter itself; the torch path refreshes the flags."""
        assert updates_dev.dtype == torch.int64 and updates_dev.is_contiguous()
        self._updates_dev = updates_dev

    def refresh_due(self):
        ud = getattr(self, "_updates_dev", None)
        if ud is not None:
            k = self.Lagrangian_multiplier_update_interval
            self.due.copy_((ud % k == 0).reshape(()).double())
            self.backup_due.copy_((ud % (k * self.backup_update_interval) == 0).reshape(()).double())


class _FusedLagrangian(torch.autograd.Function):
    """``nlbac_lagrangian_fwd / _bwd``: the whole bookkeeping + loss as one launch per direction."""

    @staticmethod
    def forward(ctx, req, st, backup):
        L = _lib.lib()
        r = req.detach().contiguous()
        m = r.numel()
        loss = torch.empty(1, dtype=torch.float32, device=r.device)
        coef = torch.empty(2 * m, dtype=torch.float32, device=r.device)
        if backup:
            lam, due = st.backup_lambda_values, st.backup_due
            rho = st.backup_augmented_term if st.separate_backup_rho else st.augmented_term
        else:
            lam, rho, due = st.lambda_values, st.augmented_term, st.due
        ud = getattr(st, "_updates_dev", None)
        interval = st.Lagrangian_multiplier_update_interval * (st.backup_update_interval if backup else 1)
        _lib.check(L.nlbac_lagrangian_fwd(
            r.data_ptr(), m, 0 if backup else 1, lam.data_ptr(), rho.data_ptr(), due.data_ptr(),
            ud.data_ptr() if ud is not None else None, interval, float(st.cost_limit),
            float(st.lambda_max), float(st.augmented_ratio), int(bool(st.use_ratio)), int(st.ratio_floor is not None),
            float(st.ratio_floor or 0.0), loss.data_ptr(), coef.data_ptr(), _lib.raw_stream()), "lagrangian_fwd")
        ctx.save_for_backward(r, coef)
        ctx.cost_limit = float(st.cost_limit)
        return loss

    @staticmethod
    def backward(ctx, g):
        r, coef = ctx.saved_tensors
        greq = torch.empty_like(r)
        _lib.check(_lib.lib().nlbac_lagrangian_bwd(r.data_ptr(), coef.data_ptr(), r.numel(), ctx.cost_limit,
                                                  g.contiguous().data_ptr(), greq.data_ptr(), _lib.raw_stream()),
                   "lagrangian_bwd")
        return greq, None, None


def _fused_lagrangian_ok(req, st):
    return (req.is_cuda and req.dtype == torch.float32 and req.numel() <= 32 and st.lambda_values.is_cuda and
            st.lambda_values.dtype == torch.float64)


def augmented_lagrangian_device(required_matrix, st):
    """``augmented_lagrangian`` (U sac_cbf_clf.py:506-530) with the multipliers, rho and the ratio kept as device
    tensors: same arithmetic (float64 scalars like the Python floats), no host synchronisation, capturable.  On CUDA
    the whole thing is ONE launch (``nlbac_lagrangian_fwd``); the torch statements below are the same arithmetic and
    remain as the path for anything the kernel does not take."""
    if _fused_lagrangian_ok(required_matrix.reshape(-1), st):
        return _FusedLagrangian.apply(required_matrix.reshape(-1), st, False).reshape(1)
    st.refresh_due()
    req = required_matrix.reshape(-1)
    c = req - st.cost_limit
    cd = c.detach().double()
    if st.use_ratio:
        ratio = torch.abs(cd[:-1].mean()) / torch.abs(cd[-1])
        if st.ratio_floor is not None:
            ratio = torch.clamp(ratio, min=st.ratio_floor)
    else:
        ratio = torch.ones((), dtype=torch.float64, device=req.device)
    new_lambda = torch.clamp(st.lambda_values + st.augmented_term * req.detach().double(), 0.01, st.lambda_max)
    st.lambda_values.copy_(torch.where(st.due > 0, new_lambda, st.lambda_values))
    st.augmented_term.copy_(torch.clamp(st.augmented_term * st.augmented_ratio, max=200.0))
    rho = st.augmented_term
    w = torch.ones_like(cd)
    w[-1] = ratio
    lin = (st.lambda_values * w).float()            # lambda_i (x ratio for the CLF row)
    quad = (rho / 2.0 * w * w).float()              # rho / 2 (x ratio^2 for the CLF row)
    return (lin * c + quad * c * c).sum().reshape(1)


def backup_augmented_lagrangian_device(backup_required_matrix, st):
    """``backup_augmented_lagrangian`` (U sac_cbf_clf.py:611-640, C :657-681, P :1022-1045) on a DeviceConstraintState."""
    if _fused_lagrangian_ok(backup_required_matrix.reshape(-1), st):
        return _FusedLagrangian.apply(backup_required_matrix.reshape(-1), st, True).reshape(1)
    st.refresh_due()
    req = backup_required_matrix.reshape(-1)
    c = req - st.cost_limit
    rho_t = st.backup_augmented_term if st.separate_backup_rho else st.augmented_term
    new_lambda = torch.clamp(st.backup_lambda_values + rho_t * req.detach().double(), 0.01, st.lambda_max)
    st.backup_lambda_values.copy_(torch.where(st.backup_due > 0, new_lambda, st.backup_lambda_values))
    rho_t.copy_(torch.clamp(rho_t * st.augmented_ratio, max=200.0))
    lin = st.backup_lambda_values.float()
    quad = (rho_t / 2.0).float()
    return (lin * c + quad * c * c).sum().reshape(1)


def _zeros_like_rows(x):
    return torch.zeros((x.shape[0], 1), dtype=torch.float32, device=x.device)


def unicycle_required_matrix(neural_ode_model, lyapunovNet, state_batch, action_batch, lyapunov_value, *, dt,
                             gamma_b, hazards_locations, hazards_radius, l_p, batch_size, solver="euler",
                             gamma_l=1.0):
    """U sac_cbf_clf.py:420-504: returns ``required_matrix`` [num_cbfs + 1, 1].  With ``lyapunovNet=None`` it is the
    CBF-only matrix of ``backup_get_policy_loss_2`` (:544-608), [num_cbfs, 1]."""
    assert state_batch.ndim == 2 and action_batch.ndim == 2, (state_batch.shape, action_batch.shape)
    ps = lookahead(state_batch, l_p)
    model_input_t = torch.cat((state_batch, action_batch), dim=-1)
    next_state_t = odeint(neural_ode_model, model_input_t, [0.0, dt], method=solver, atol=1e-7, rtol=1e-5)[-1]
    ps_next = lookahead(next_state_t[:, :3], l_p)
    desc = make_constraint_desc("unicycle", dt=dt, gamma_b=gamma_b, gamma_l=gamma_l, batch_size=batch_size,
                                collision_radius=1.05 * hazards_radius, hazards=hazards_locations)
    if lyapunovNet is None:   # zero CLF row, dropped
        z = _zeros_like_rows(ps)
        return constraint_sums(desc, ps, ps_next, None, z, z)[:-1]
    lyapunov_value_next = lyapunovNet(ps_next)
    return constraint_sums(desc, ps, ps_next, None, lyapunov_value.detach(), lyapunov_value_next)


def cars_required_matrix(neural_ode_model, lyapunovNet, policy, dynamics_model, state_batch, action_batch,
                         lyapunov_value, t_batch, next_time_batch, *, dt, gamma_b, batch_size, solver="euler",
                         gamma_l=0.15, eps_next=None):
    """C sac_cbf_clf.py:424-523: two chained NODE predictions with the policy re-sampled on the detached
    predicted observation in between; returns ``required_matrix`` [3, 1].  With ``lyapunovNet=None`` (and ``policy`` =
    the backup policy) it is the CBF-only matrix of ``backup_get_policy_loss_2`` (:557-645), [2, 1]."""
    model_input_t = torch.cat((state_batch, action_batch, t_batch), dim=-1)
    next_state_pred = odeint(neural_ode_model, model_input_t, [0.0, dt], method=solver, atol=1e-7, rtol=1e-5)[-1][:, :10]
    next_obs_pred = dynamics_model.get_obs(next_state_pred.detach())
    pi_next, _, _ = policy.sample(next_obs_pred, eps_next)
    pi_next = pi_next.detach()
    next_model_input_t = torch.cat((next_state_pred, pi_next, next_time_batch), dim=-1)
    next_next_state_pred = odeint(neural_ode_model, next_model_input_t, [0.0, dt], method=solver, atol=1e-7,
                                  rtol=1e-5)[-1][:, :10]
    desc = make_constraint_desc("cars", dt=dt, gamma_b=gamma_b, gamma_l=gamma_l, batch_size=batch_size,
                                lya_inv_dt=1.0, collision_radius=4.5)
    if lyapunovNet is None:
        z = _zeros_like_rows(state_batch)
        return constraint_sums(desc, state_batch, next_state_pred, next_next_state_pred, z, z)[:-1]
    lyapunov_value_next = lyapunovNet(next_state_pred[:, 4:8])
    return constraint_sums(desc, state_batch, next_state_pred, next_next_state_pred, lyapunov_value.detach(),
                           lyapunov_value_next)


def barrier_unicycle_required_matrix(neural_ode_model, lyapunovNet, BarrierNet, policy, dynamics_model, obs_batch,
                                     state_batch, action_batch, lyapunov_value, *, dt, gamma_b, l_p, batch_size,
                                     solver="euler", gamma_l=1.0, eps_next=None):
    """BU sac_cbf_clf.py:368-446: learned barrier certificate instead of closed-form CBFs; ``required_matrix`` [2, 1]."""
    model_input_t = torch.cat((state_batch, action_batch), dim=-1)
    next_state_t = odeint(neural_ode_model, model_input_t, [0.0, dt], method=solver, atol=1e-7, rtol=1e-5)[-1]
    predicted = next_state_t[:, :3]
    ps_next = lookahead(predicted, l_p)
    lyapunov_value_next = lyapunovNet(ps_next)
    barrier_value = BarrierNet(obs_batch, action_batch).detach()
    next_obs = dynamics_model.get_obs(predicted)
    pi_next, _, _ = policy.sample(next_obs.detach(), eps_next)
    barrier_next = BarrierNet(next_obs, pi_next.detach())
    desc = make_constraint_desc("barrier", dt=dt, gamma_b=gamma_b, gamma_l=gamma_l, batch_size=batch_size)
    return constraint_sums(desc, barrier_value, barrier_next, None, lyapunov_value.detach(), lyapunov_value_next)


def barrier_pvtol_required_matrix(neural_ode_model, lyapunovNet, BarrierNet, policy, dynamics_model, obs_batch,
                                  state_batch, action_batch, lyapunov_value, state_dynamics_batch=None, *, dt, gamma_b,
                                  safety_operator_follow, batch_size, solver="euler", gamma_l=0.1, eps_next=None):
    """BP sac_cbf_clf.py:373-438: PVTOL with a learned barrier certificate; ``required_matrix`` [2, 1]."""
    if state_dynamics_batch is None:
        state_dynamics_batch = state_batch[:, :6]
    nxt = odeint(neural_ode_model, torch.cat((state_dynamics_batch, action_batch), dim=-1), [0.0, dt], method=solver,
                 atol=1e-7, rtol=1e-5)[-1][:, :6]
    op = state_batch[:, 6] + safety_operator_follow * (nxt[:, 0] - state_batch[:, 6])
    next_obs = dynamics_model.get_obs(torch.cat((nxt, op[:, None]), dim=1))
    lyapunov_value_next = lyapunovNet(next_obs)
    pi_next, _, _ = policy.sample(next_obs.detach(), eps_next)
    barrier_value = BarrierNet(obs_batch, action_batch).detach()
    barrier_next = BarrierNet(next_obs, pi_next.detach())
    desc = make_constraint_desc("barrier", dt=dt, gamma_b=gamma_b, gamma_l=gamma_l, batch_size=batch_size,
                                lya_inv_dt=1.0)
    return constraint_sums(desc, barrier_value, barrier_next, None, lyapunov_value.detach(), lyapunov_value_next)


def pvtol_required_matrix(neural_ode_model, lyapunovNet, policy, dynamics_model, state_batch, action_batch,
                          lyapunov_value, state_dynamics_batch=None, *, dt, gamma_b, hazard_locations, hazards_radius,
                          safety_operator_follow, operator_dist, y_max, y_min, batch_size, solver="euler",
                          gamma_l=0.1, eps_next=None, eps_next_next=None):
    """P sac_cbf_clf.py:424-697: three chained NODE predictions of the six dynamic states (the seventh state, the
    safety operator's x, follows ``op + follow * (x_next - op)``, :461-464), the policy re-sampled on the detached
    predicted observation between steps (:483-489, :520-527), then 5 hazard + 4 operator / altitude third-order CBF
    rows and the CLF row in ONE fused kernel; returns ``required_matrix`` [10, 1].  With ``lyapunovNet=None`` (and
    ``policy`` = the backup policy) it is the CBF-only matrix of ``backup_get_policy_loss_2`` (:740-1010), [9, 1]."""
    assert state_batch.ndim == 2 and action_batch.ndim == 2, (state_batch.shape, action_batch.shape)
    if state_dynamics_batch is None:
        state_dynamics_batch = state_batch[:, :6]

    def step(s7_prev, dyn_prev, u):
        nxt = odeint(neural_ode_model, torch.cat((dyn_prev, u), dim=-1), [0.0, dt], method=solver, atol=1e-7,
                     rtol=1e-5)[-1][:, :6]
        op = s7_prev[:, 6] + safety_operator_follow * (nxt[:, 0] - s7_prev[:, 6])
        return torch.cat((nxt, op[:, None]), dim=1), nxt

    s1, d1 = step(state_batch, state_dynamics_batch, action_batch)
    next_obs = dynamics_model.get_obs(s1)
    pi_next = policy.sample(next_obs.detach(), eps_next)[0].detach()
    s2, d2 = step(s1, d1, pi_next)
    pi_next_next = policy.sample(dynamics_model.get_obs(s2.detach()), eps_next_next)[0].detach()
    s3, _ = step(s2, d2, pi_next_next)
    desc = make_constraint_desc("pvtol", dt=dt, gamma_b=gamma_b, gamma_l=gamma_l, batch_size=batch_size,
                                lya_inv_dt=1.0, collision_radius=1.2 * hazards_radius, hazards=hazard_locations,
                                operator_dist=0.9 * operator_dist, y_max=y_max, y_min=y_min, delta_y=10.0)
    if lyapunovNet is None:
        z = _zeros_like_rows(state_batch)
        return constraint_sums(desc, state_batch, s1, s2, z, z, e=s3)[:-1]
    lyapunov_value_next = lyapunovNet(next_obs)
    return constraint_sums(desc, state_batch, s1, s2, lyapunov_value.detach(), lyapunov_value_next, e=s3)


def chained_rollout(neural_ode_model, policy, state_batch, action_batch, horizon, *, dt, eps=None, solver="euler",
                    obs_fn=None, const_cols=None):
    """Multi-step NODE rollout with the policy re-sampled on the DETACHED predicted observation between steps -- the
    pattern of C sac_cbf_clf.py:437-459 (2 steps) and P :459-552 (3 steps) at an arbitrary horizon (BASELINE config 4:
    the Quadrotor-shaped model, horizon 10, ``eps[horizon - 1, B, n_u]`` injected; README.md:190-192).
    One fused solve per step; gradients flow through the chain of predicted states to ``action_batch`` and the NODE
    parameters, not through the re-sampled actions.  ``obs_fn`` maps a predicted state to the policy's input (identity
    by default); ``const_cols`` are extra constant columns of y0 after the action (the cars' time feature), one
    tensor per step.  Returns the predicted states ``[horizon, B, n_s]``."""
    n_s = state_batch.shape[1]
    x, u, out = state_batch, action_batch, []
    for k in range(horizon):
        cols = (x, u) if const_cols is None else (x, u, const_cols[k])
        x = odeint(neural_ode_model, torch.cat(cols, dim=-1), [0.0, dt], method=solver, atol=1e-7, rtol=1e-5)[-1][:, :n_s]
        out.append(x)
        if k + 1 < horizon:
            obs = x.detach() if obs_fn is None else obs_fn(x.detach())
            u = policy.sample(obs, None if eps is None else eps[k])[0].detach()
    return torch.stack(out)
