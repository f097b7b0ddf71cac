"""Fused element-wise / reduction tail of the update step over the C ABI.

* ``constraint_sums``  — CBF/CLF rows -> ``where(>0)`` -> batch sum / ``batch_size`` (U sac_cbf_clf.py:472-504,
  C :474-523, BU :424-460), one kernel forward, one backward.
* ``lookahead``        — ``p(x) = [x, y] + l_p [cos th, sin th]`` (U sac_cbf_clf.py:429-437).
* ``DynamicsModel``    — ``get_state`` / ``get_obs`` with the reference's signatures (U dynamics.py:27-69,
  C :25-96, P :25-153, BU :73-198); tensors never leave the device.
* ``FusedAdam`` / ``soft_update`` — torch.optim.Adam.step on flat tensors and Polyak averaging (U utils.py:75-79).
"""
import ctypes

import numpy as np
import torch

from . import _lib

CONSTRAINT_KINDS = {"unicycle": 0, "cars": 1, "barrier": 2, "pvtol": 3}
ENVS = {"Unicycle": 0, "SimulatedCars": 1, "Pvtol": 2}
_OBS_DIMS = {0: (7, 3), 1: (10, 10), 2: (11, 7)}


def _stream():
    return _lib.raw_stream()


def _cuda32(*ts):
    out = []
    for t in ts:
        if t is None:
            out.append(None)
            continue
        if not t.is_cuda:
            raise RuntimeError("nlbac_b200: tensors must be CUDA tensors (no CPU fallback)")
        out.append(t.detach().contiguous().float())
    return out


def _p(t):
    return None if t is None else t.data_ptr()


def make_constraint_desc(kind, *, dt, gamma_b, gamma_l, batch_size, lya_inv_dt=None, collision_radius=0.0,
                         hazards=None, operator_dist=0.0, y_max=0.0, y_min=0.0, delta_y=10.0):
    d = _lib.ConstraintDesc()
    d.kind = CONSTRAINT_KINDS[kind]
    hz = np.zeros((0, 2), dtype=np.float32) if hazards is None else np.asarray(hazards, dtype=np.float32).reshape(-1, 2)
    if len(hz) > _lib.MAX_HAZARDS:
        raise ValueError("at most %d hazards" % _lib.MAX_HAZARDS)
    d.n_hazards = len(hz) if kind in ("unicycle", "pvtol") else 0
    d.m = {"unicycle": d.n_hazards + 1, "cars": 3, "barrier": 2, "pvtol": d.n_hazards + 5}[kind]
    d.operator_dist, d.y_max, d.y_min, d.delta_y = float(operator_dist), float(y_max), float(y_min), float(delta_y)
    d.dt, d.gamma_b, d.gamma_l = float(dt), float(gamma_b), float(gamma_l)
    d.lya_inv_dt = float(1.0 / dt if lya_inv_dt is None else lya_inv_dt)
    d.batch_size = float(batch_size)
    d.collision_radius = float(collision_radius)
    for i, v in enumerate(hz.reshape(-1)):
        d.hazards[i] = float(v)
    return d


class _ConstraintSums(torch.autograd.Function):
    @staticmethod
    def forward(ctx, desc, a, b, c, e, V, Vnext):
        L = _lib.lib()
        a_, b_, c_, e_, V_, Vn_ = _cuda32(a, b, c, e, V, Vnext)
        B = V_.shape[0]
        req = torch.empty((desc.m, 1), dtype=torch.float32, device=V_.device)
        _lib.check(L.nlbac_constraints4_fwd(ctypes.byref(desc), _p(a_), _p(b_), _p(c_), _p(e_), _p(V_), _p(Vn_), B,
                                            req.data_ptr(), _stream()), "constraints_fwd")
        ctx.desc = desc
        ctx.has = (c_ is not None, e_ is not None)
        ctx.save_for_backward(a_, b_, V_, Vn_, *[t for t in (c_, e_) if t is not None])
        ctx.shapes = (b.shape, None if c is None else c.shape, None if e is None else e.shape, Vnext.shape)
        return req

    @staticmethod
    def backward(ctx, g):
        L = _lib.lib()
        a_, b_, V_, Vn_, *rest = ctx.saved_tensors
        c_ = rest.pop(0) if ctx.has[0] else None
        e_ = rest.pop(0) if ctx.has[1] else None
        B = V_.shape[0]
        gb = torch.empty_like(b_)
        gc = torch.empty_like(c_) if c_ is not None else None
        ge = torch.empty_like(e_) if e_ is not None else None
        gV = torch.empty_like(Vn_)
        gq = g.contiguous().float()
        _lib.check(L.nlbac_constraints4_bwd(ctypes.byref(ctx.desc), _p(a_), _p(b_), _p(c_), _p(e_), _p(V_), _p(Vn_), B,
                                            gq.data_ptr(), gb.data_ptr(), _p(gc), _p(ge), gV.data_ptr(), _stream()),
                   "constraints_bwd")
        sb, sc, se, sv = ctx.shapes
        return (None, None, gb.view(sb), (gc.view(sc) if gc is not None else None),
                (ge.view(se) if ge is not None else None), None, gV.view(sv))


def constraint_sums(desc, a, b, c, V, Vnext, e=None):
    """``required_matrix`` [m, 1] of the reference: filtered constraint rows summed over the batch and divided
    by ``batch_size``.  Gradients flow to ``b`` (x_{t+1} quantities), ``c`` (x_{t+2}), ``e`` (x_{t+3}) and ``Vnext``."""
    return _ConstraintSums.apply(desc, a, b, c, e, V, Vnext)


class _Lookahead(torch.autograd.Function):
    @staticmethod
    def forward(ctx, state, l_p):
        L = _lib.lib()
        (s,) = _cuda32(state)
        B, n_s = s.shape
        p = torch.empty((B, 2), dtype=torch.float32, device=s.device)
        _lib.check(L.nlbac_lookahead_fwd(s.data_ptr(), n_s, float(l_p), B, p.data_ptr(), _stream()), "lookahead_fwd")
        ctx.save_for_backward(s)
        ctx.l_p = float(l_p)
        return p

    @staticmethod
    def backward(ctx, gp):
        L = _lib.lib()
        (s,) = ctx.saved_tensors
        B, n_s = s.shape
        gs = torch.empty_like(s)
        g = gp.contiguous().float()
        _lib.check(L.nlbac_lookahead_bwd(s.data_ptr(), n_s, ctx.l_p, B, g.data_ptr(), gs.data_ptr(), _stream()),
                   "lookahead_bwd")
        return gs, None


def lookahead(state, l_p):
    return _Lookahead.apply(state, l_p)


class _GetObs(torch.autograd.Function):
    @staticmethod
    def forward(ctx, state, env):
        L = _lib.lib()
        (s,) = _cuda32(state)
        B = s.shape[0]
        obs = torch.empty((B, _OBS_DIMS[env][0]), dtype=torch.float32, device=s.device)
        _lib.check(L.nlbac_get_obs_fwd(env, s.data_ptr(), B, obs.data_ptr(), _stream()), "get_obs_fwd")
        ctx.save_for_backward(s)
        ctx.env = env
        return obs

    @staticmethod
    def backward(ctx, g):
        L = _lib.lib()
        (s,) = ctx.saved_tensors
        gs = torch.empty_like(s)
        gc = g.contiguous().float()
        _lib.check(L.nlbac_get_obs_bwd(ctx.env, s.data_ptr(), s.shape[0], gc.data_ptr(), gs.data_ptr(), _stream()),
                   "get_obs_bwd")
        return gs, None


class DynamicsModel:
    """``DynamicsModel(env, args)`` with ``get_state`` / ``get_obs`` like the reference (U dynamics.py:10-76).

    ``env`` needs only ``dynamics_mode`` ('Unicycle' | 'SimulatedCars' | 'Pvtol').  Tensor inputs stay on the
    GPU (the reference round-trips through float64 numpy, App. D.1; the kernel does the same arithmetic in
    float64 registers); ndarray inputs follow the reference's numpy path on the host, as in ``select_action``.
    """

    def __init__(self, env, args=None):
        self.env = env
        if env.dynamics_mode not in ENVS:
            raise Exception('Unknown dynamics')
        self.kind = ENVS[env.dynamics_mode]

    def seed(self, s):
        torch.manual_seed(s)

    def get_state(self, obs):
        if not torch.is_tensor(obs):
            return self._get_state_numpy(np.asarray(obs))
        expand = obs.ndim == 1
        o = obs.unsqueeze(0) if expand else obs
        (oc,) = _cuda32(o)
        n_o, n_s = _OBS_DIMS[self.kind]
        if oc.shape[1] != n_o:
            raise ValueError("observation width %d != %d" % (oc.shape[1], n_o))
        B = oc.shape[0]
        L = _lib.lib()
        state = torch.empty((B, n_s), dtype=torch.float32, device=oc.device)
        dyn = torch.empty((B, 6), dtype=torch.float32, device=oc.device) if self.kind == 2 else None
        _lib.check(L.nlbac_get_state(self.kind, oc.data_ptr(), B, state.data_ptr(), _p(dyn), _stream()), "get_state")
        state = state.to(obs.dtype)
        if self.kind == 2:
            dyn = dyn.to(obs.dtype)
            return (state.squeeze(0), dyn.squeeze(0)) if expand else (state, dyn)
        return state.squeeze(0) if expand else state

    def _get_state_numpy(self, obs):
        expand = obs.ndim == 1
        o = obs[None] if expand else obs
        if self.kind == 0:
            st = np.stack([o[:, 0], o[:, 1], np.arctan2(o[:, 3], o[:, 2])], 1)
        elif self.kind == 1:
            st = np.copy(o)
            st[:, ::2] *= 100.0
            st[:, 1::2] *= 30.0
        else:
            st = np.stack([o[:, 0], o[:, 1], np.arctan2(o[:, 3], o[:, 2]), o[:, 4], o[:, 5], o[:, 6], o[:, 7]], 1)
            dyn = st[:, :6]
            return (st[0], dyn[0]) if expand else (st, dyn)
        return st[0] if expand else st

    def get_obs(self, state_batch, device=None):
        """``state_batch`` is ``[B, n_s, 1]`` like the reference (or ``[B, n_s]``); returns the same rank."""
        col = state_batch.ndim == 3
        s = state_batch.squeeze(-1) if col else state_batch
        obs = _GetObs.apply(s, self.kind)
        return obs.unsqueeze(-1) if col else obs


class FusedAdam(torch.optim.Optimizer):
    """torch.optim.Adam (default betas/eps, no weight decay) with the update run by ``nlbac_adam_step_multi``: one
    launch per parameter group.  ``capturable=True`` keeps the step counter on the device
    (``nlbac_adam_step_multi_dev``) so that ``step()`` can be captured in a CUDA graph and replayed."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, capturable=False):
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps))
        self.capturable = bool(capturable)
        self._step_dev = {}

    @torch.no_grad()
    def step(self, closure=None):
        L = _lib.lib()
        for group in self.param_groups:
            b1, b2 = group["betas"]
            if self.capturable:
                self._step_capturable(L, group, b1, b2)
                continue
            by_step = {}
            for p in group["params"]:
                if p.grad is None:
                    continue
                if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous():
                    raise RuntimeError("FusedAdam: parameters must be contiguous float32 CUDA tensors")
                st = self.state[p]
                if not st:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p)
                    st["exp_avg_sq"] = torch.zeros_like(p)
                st["step"] += 1
                by_step.setdefault(st["step"], []).append((p, p.grad.contiguous(), st))
            # one launch per (group, step count): normally ONE launch for the whole network
            for step, items in by_step.items():
                for lo in range(0, len(items), _lib.MAX_TENSORS):
                    chunk = items[lo:lo + _lib.MAX_TENSORS]
                    n = len(chunk)
                    arr = lambda xs: (ctypes.c_void_p * n)(*xs)  # noqa: E731
                    _lib.check(L.nlbac_adam_step_multi(
                        n, arr([p.data_ptr() for p, _, _ in chunk]), arr([g.data_ptr() for _, g, _ in chunk]),
                        arr([s["exp_avg"].data_ptr() for _, _, s in chunk]),
                        arr([s["exp_avg_sq"].data_ptr() for _, _, s in chunk]),
                        (ctypes.c_int64 * n)(*[p.numel() for p, _, _ in chunk]), step, group["lr"], b1, b2,
                        group["eps"], _stream()), "adam_step_multi")
                    for p, _, _ in chunk:
                        torch.autograd.graph.increment_version(p)   # written through raw pointers: invalidate caches


def _fused_adam_step_capturable(self, L, group, b1, b2):
    items = []
    for p in group["params"]:
        if p.grad is None:
            continue
        if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous():
            raise RuntimeError("FusedAdam: parameters must be contiguous float32 CUDA tensors")
        st = self.state[p]
        if not st:
            st["exp_avg"] = torch.zeros_like(p)
            st["exp_avg_sq"] = torch.zeros_like(p)
        items.append((p, p.grad.contiguous(), st))
    if not items:
        return
    key = id(group)
    if key not in self._step_dev:
        self._step_dev[key] = torch.zeros(2, dtype=torch.int32, device=items[0][0].device)   # step, ticket word
    step_dev = self._step_dev[key]
    for lo in range(0, len(items), _lib.MAX_TENSORS):
        chunk = items[lo:lo + _lib.MAX_TENSORS]
        n = len(chunk)
        arr = lambda xs: (ctypes.c_void_p * n)(*xs)  # noqa: E731
        _lib.check(L.nlbac_adam_step_multi_dev(
            n, arr([p.data_ptr() for p, _, _ in chunk]), arr([g.data_ptr() for _, g, _ in chunk]),
            arr([s["exp_avg"].data_ptr() for _, _, s in chunk]), arr([s["exp_avg_sq"].data_ptr() for _, _, s in chunk]),
            (ctypes.c_int64 * n)(*[p.numel() for p, _, _ in chunk]), step_dev.data_ptr(), 1 if lo == 0 else 0,
            group["lr"], b1, b2, group["eps"], _stream()), "adam_step_multi_dev")
        for p, _, _ in chunk:
            torch.autograd.graph.increment_version(p)


FusedAdam._step_capturable = _fused_adam_step_capturable


@torch.no_grad()
def soft_update(target, source, tau):
    """``utils.soft_update(target, source, tau)`` (U utils.py:75-79): one launch for the whole network."""
    L = _lib.lib()
    pairs = list(zip(target.parameters(), source.parameters()))
    for lo in range(0, len(pairs), _lib.MAX_TENSORS):
        chunk = pairs[lo:lo + _lib.MAX_TENSORS]
        n = len(chunk)
        for tp, sp in chunk:
            if not (tp.is_cuda and sp.is_cuda and tp.is_contiguous() and sp.is_contiguous()
                    and tp.dtype == sp.dtype == torch.float32 and tp.numel() == sp.numel()):
                raise RuntimeError("soft_update: parameters must be matching contiguous float32 CUDA tensors")
        _lib.check(L.nlbac_soft_update_multi(
            n, (ctypes.c_void_p * n)(*[tp.data_ptr() for tp, _ in chunk]),
            (ctypes.c_void_p * n)(*[sp.data_ptr() for _, sp in chunk]),
            (ctypes.c_int64 * n)(*[tp.numel() for tp, _ in chunk]), float(tau), _stream()), "soft_update_multi")
        for tp, _ in chunk:
            torch.autograd.graph.increment_version(tp)


@torch.no_grad()
def hard_update(target, source):
    for tp, sp in zip(target.parameters(), source.parameters()):
        tp.copy_(sp)


# ------------------------------------------------------------------------------------------------
# scalar glue of update_parameters as single launches
# ------------------------------------------------------------------------------------------------
def _f32c(x):
    if not (x.is_cuda and x.dtype == torch.float32):
        raise RuntimeError("float32 CUDA tensor expected")
    return x.detach().contiguous()


def td_target(reward, mask, gamma, q1, q2=None, logp=None, alpha=None):
    """``reward + mask * gamma * (min(q1, q2) - alpha * logp)`` (U sac_cbf_clf.py:234-236) in one launch; with ``q2`` /
    ``logp`` omitted the Lyapunov / barrier target ``reward + mask * gamma * q1`` (:237-238).  No gradient (the
    reference computes it under ``no_grad``).  ``alpha``: float32 CUDA tensor [1]."""
    r, m, a = _f32c(reward), _f32c(mask), _f32c(q1)
    b2 = _f32c(q2) if q2 is not None else None
    lp = _f32c(logp) if logp is not None else None
    al = _f32c(alpha) if logp is not None else None
    out = torch.empty_like(a)
    _lib.check(_lib.lib().nlbac_td_target(a.data_ptr(), b2.data_ptr() if b2 is not None else None,
                                          lp.data_ptr() if lp is not None else None,
                                          al.data_ptr() if al is not None else None, r.data_ptr(), m.data_ptr(),
                                          float(gamma), a.numel(), out.data_ptr(), _stream()), "td_target")
    return out


class _MseMean(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, target):
        p, t = _f32c(pred), _f32c(target)
        assert p.shape == t.shape
        loss = torch.empty((), dtype=torch.float32, device=p.device)
        _lib.check(_lib.lib().nlbac_mse_mean_fwd(p.data_ptr(), t.data_ptr(), p.numel(), loss.data_ptr(), _stream()),
                   "mse_mean_fwd")
        ctx.save_for_backward(p, t)
        return loss

    @staticmethod
    def backward(ctx, g):
        p, t = ctx.saved_tensors
        gp = torch.empty_like(p)
        _lib.check(_lib.lib().nlbac_mse_mean_bwd(p.data_ptr(), t.data_ptr(), p.numel(), g.contiguous().data_ptr(),
                                                 gp.data_ptr(), _stream()), "mse_mean_bwd")
        return gp, None


def mse_mean(pred, target):
    """``F.mse_loss(pred, target)`` (mean reduction; the target is a constant) as one launch per direction."""
    return _MseMean.apply(pred, target)


class _SacActorTerm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, log_pi, q1, q2, alpha):
        lp, a, b, al = _f32c(log_pi), _f32c(q1), _f32c(q2), _f32c(alpha)
        loss = torch.empty((), dtype=torch.float32, device=lp.device)
        _lib.check(_lib.lib().nlbac_sac_actor_term_fwd(lp.data_ptr(), a.data_ptr(), b.data_ptr(), al.data_ptr(),
                                                       lp.numel(), loss.data_ptr(), _stream()), "sac_actor_term_fwd")
        ctx.save_for_backward(a, b, al)
        return loss

    @staticmethod
    def backward(ctx, g):
        a, b, al = ctx.saved_tensors
        glp, ga, gb = torch.empty_like(a), torch.empty_like(a), torch.empty_like(b)
        _lib.check(_lib.lib().nlbac_sac_actor_term_bwd(a.data_ptr(), b.data_ptr(), al.data_ptr(), a.numel(),
                                                       g.contiguous().data_ptr(), glp.data_ptr(), ga.data_ptr(),
                                                       gb.data_ptr(), _stream()), "sac_actor_term_bwd")
        return glp, ga, gb, None


def sac_actor_term(log_pi, q1, q2, alpha):
    """``(alpha * log_pi - torch.min(q1, q2)).mean()`` (U sac_cbf_clf.py:270) as one launch per direction; ``alpha`` is a
    constant float32 CUDA tensor [1] (a Python float in the reference)."""
    return _SacActorTerm.apply(log_pi, q1, q2, alpha)
