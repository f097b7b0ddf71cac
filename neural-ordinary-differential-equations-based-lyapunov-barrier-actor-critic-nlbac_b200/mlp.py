"""Host mirror of the reference's actor / critic / Lyapunov / barrier networks.

Same class names, constructor signatures, attribute names (``linear1..6``, ``mean_linear``,
``log_std_linear``) and state-dict keys as U ``sac_cbf_clf/model.py:37-133`` and BU ``model.py:67-86``;
xavier-uniform weights and zero biases (``weights_init_``, U model.py:14-17).  ``forward`` runs the fused
CUDA MLP (one launch forward; rows kernel + batched A^T B + column sums backward) through the C ABI;
there is no eager fallback.
"""
import ctypes

import torch
import torch.nn as nn

from . import _lib

LOG_SIG_MAX = 2
LOG_SIG_MIN = -20
epsilon = 1e-6


def weights_init_(m):
    if isinstance(m, nn.Linear):
        torch.nn.init.xavier_uniform_(m.weight, gain=1)
        torch.nn.init.constant_(m.bias, 0)


def _ptrs(tensors):
    return (ctypes.c_void_p * len(tensors))(*[t.data_ptr() for t in tensors])


def _stream():
    return _lib.raw_stream()


class _MLP3(torch.autograd.Function):
    """out = W3 relu(W2 relu(W1 x + b1) + b2) + b3, fused."""

    @staticmethod
    def forward(ctx, x, W1, b1, W2, b2, W3, b3):
        if not x.is_cuda:
            raise RuntimeError("nlbac_b200 MLP: input must be a CUDA tensor (no CPU fallback)")
        L = _lib.lib()
        params = [p.detach().contiguous() for p in (W1, b1, W2, b2, W3, b3)]
        xc = x.detach().contiguous().float()
        B = xc.shape[0]
        desc = _lib.MlpDesc(W1.shape[1], W1.shape[0], W3.shape[0])
        need = any(ctx.needs_input_grad)
        out = torch.empty((B, desc.n_out), dtype=torch.float32, device=x.device)
        act = torch.empty((2, B, desc.hidden), dtype=torch.float32, device=x.device) if need else None
        _lib.check(L.nlbac_mlp_fwd(ctypes.byref(desc), _ptrs(params), xc.data_ptr(), B, out.data_ptr(),
                                   act.data_ptr() if need else None, _stream()), "mlp_fwd")
        if need:
            ctx.save_for_backward(xc, act, *params)
            ctx.desc = desc
        return out

    @staticmethod
    def backward(ctx, g):
        L = _lib.lib()
        xc, act, *params = ctx.saved_tensors
        desc = ctx.desc
        B = xc.shape[0]
        dev = xc.device
        need_x = ctx.needs_input_grad[0]
        need_p = any(ctx.needs_input_grad[1:])
        gx = torch.empty_like(xc) if need_x else None
        gp = [torch.empty_like(p) for p in params] if need_p else None
        ws = torch.empty((2, B, desc.hidden), dtype=torch.float32, device=dev)
        gc = g.contiguous().float()
        _lib.check(L.nlbac_mlp_bwd(ctypes.byref(desc), _ptrs(params), xc.data_ptr(), B, act.data_ptr(), gc.data_ptr(),
                                   gx.data_ptr() if need_x else None, _ptrs(gp) if need_p else None, ws.data_ptr(),
                                   _stream()), "mlp_bwd")
        return (gx, *(gp if need_p else [None] * 6))


def mlp3(x, l1, l2, l3):
    return _MLP3.apply(x, l1.weight, l1.bias, l2.weight, l2.bias, l3.weight, l3.bias)


class _PolicyHead(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ml, eps, scale, bias):
        L = _lib.lib()
        B, two_nu = ml.shape
        nu = two_nu // 2
        mlc, epc = ml.detach().contiguous(), eps.detach().contiguous().float()
        action = torch.empty((B, nu), dtype=torch.float32, device=ml.device)
        logp = torch.empty((B, 1), dtype=torch.float32, device=ml.device)
        mean = torch.empty((B, nu), dtype=torch.float32, device=ml.device)
        _lib.check(L.nlbac_policy_head_fwd(mlc.data_ptr(), epc.data_ptr(), scale.data_ptr(), bias.data_ptr(), B, nu,
                                           action.data_ptr(), logp.data_ptr(), mean.data_ptr(), _stream()),
                   "policy_head_fwd")
        ctx.save_for_backward(mlc, epc, scale)
        return action, logp, mean

    @staticmethod
    def backward(ctx, ga, gl, gm):
        L = _lib.lib()
        mlc, epc, scale = ctx.saved_tensors
        B, two_nu = mlc.shape
        g = torch.empty_like(mlc)
        ga, gl, gm = [None if t is None else t.contiguous().float() for t in (ga, gl, gm)]
        _lib.check(L.nlbac_policy_head_bwd(mlc.data_ptr(), epc.data_ptr(), scale.data_ptr(), B, two_nu // 2,
                                           ga.data_ptr() if ga is not None else None,
                                           gl.data_ptr() if gl is not None else None,
                                           gm.data_ptr() if gm is not None else None, g.data_ptr(), _stream()),
                   "policy_head_bwd")
        return g, None, None, None


_TWIN_STREAMS = {}


def _twin_side_stream(cur):
    """One side stream per caller stream (two branches of a captured update may evaluate the same critic at once)."""
    key = (cur.device_index, cur.cuda_stream)
    if key not in _TWIN_STREAMS:
        _TWIN_STREAMS[key] = torch.cuda.Stream(device=cur.device)
    return _TWIN_STREAMS[key]


class QNetwork(nn.Module):
    """Twin critics (U model.py:37-64): ``forward(state, action) -> (q1, q2)``."""

    def __init__(self, num_inputs, num_actions, hidden_dim):
        super().__init__()
        self.linear1 = nn.Linear(num_inputs + num_actions, hidden_dim)
        self.linear2 = nn.Linear(hidden_dim, hidden_dim)
        self.linear3 = nn.Linear(hidden_dim, 1)
        self.linear4 = nn.Linear(num_inputs + num_actions, hidden_dim)
        self.linear5 = nn.Linear(hidden_dim, hidden_dim)
        self.linear6 = nn.Linear(hidden_dim, 1)
        self.apply(weights_init_)

    # The two critics are independent: with ``twin_streams`` the second one runs on a side stream forked from the
    # caller's (its backward follows it there: autograd replays each node on its forward stream).  At B = 128 / 256
    # one critic fills < half of the SMs; GraphedUpdate switches this on, where the fork costs nothing at replay.
    twin_streams = False

    def forward(self, state, action):
        xu = torch.cat([state, action], 1)
        if not (self.twin_streams and xu.is_cuda):
            return mlp3(xu, self.linear1, self.linear2, self.linear3), mlp3(xu, self.linear4, self.linear5, self.linear6)
        cur = torch.cuda.current_stream()
        side = _twin_side_stream(cur)
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            q2 = mlp3(xu, self.linear4, self.linear5, self.linear6)
        q1 = mlp3(xu, self.linear1, self.linear2, self.linear3)
        cur.wait_stream(side)
        return q1, q2


class LyaNetwork(nn.Module):
    """U model.py:67-83."""

    def __init__(self, num_inputs, hidden_dim):
        super().__init__()
        self.linear1 = nn.Linear(num_inputs, hidden_dim)
        self.linear2 = nn.Linear(hidden_dim, hidden_dim)
        self.linear3 = nn.Linear(hidden_dim, 1)
        self.apply(weights_init_)

    def forward(self, state):
        return mlp3(state, self.linear1, self.linear2, self.linear3)


class BarrierNetwork(nn.Module):
    """BU model.py:67-86, BP model.py:68-87."""

    def __init__(self, num_inputs, num_actions, hidden_dim):
        super().__init__()
        self.linear1 = nn.Linear(num_inputs + num_actions, hidden_dim)
        self.linear2 = nn.Linear(hidden_dim, hidden_dim)
        self.linear3 = nn.Linear(hidden_dim, 1)
        self.apply(weights_init_)

    def forward(self, state, action):
        return mlp3(torch.cat([state, action], 1), self.linear1, self.linear2, self.linear3)


class GaussianPolicy(nn.Module):
    """U model.py:86-133.  ``sample(state)`` draws eps from torch's global RNG like the reference's
    ``Normal.rsample``; ``sample(state, eps=...)`` injects it (parity tests, fused rollouts)."""

    def __init__(self, num_inputs, num_actions, hidden_dim, action_space=None):
        super().__init__()
        self.linear1 = nn.Linear(num_inputs, hidden_dim)
        self.linear2 = nn.Linear(hidden_dim, hidden_dim)
        self.mean_linear = nn.Linear(hidden_dim, num_actions)
        self.log_std_linear = nn.Linear(hidden_dim, num_actions)
        self.apply(weights_init_)
        if action_space is None:
            self.action_scale = torch.ones(num_actions)
            self.action_bias = torch.zeros(num_actions)
        else:
            self.action_scale = torch.FloatTensor((action_space.high - action_space.low) / 2.)
            self.action_bias = torch.FloatTensor((action_space.high + action_space.low) / 2.)

    def _mean_logstd(self, state):
        W3 = torch.cat([self.mean_linear.weight, self.log_std_linear.weight], 0)
        b3 = torch.cat([self.mean_linear.bias, self.log_std_linear.bias], 0)
        return _MLP3.apply(state, self.linear1.weight, self.linear1.bias, self.linear2.weight, self.linear2.bias, W3, b3)

    def forward(self, state):
        ml = self._mean_logstd(state)
        nu = ml.shape[1] // 2
        return ml[:, :nu], torch.clamp(ml[:, nu:], min=LOG_SIG_MIN, max=LOG_SIG_MAX)

    def sample(self, state, eps=None):
        ml = self._mean_logstd(state)
        if eps is None:
            eps = torch.randn((ml.shape[0], ml.shape[1] // 2), dtype=ml.dtype, device=ml.device)
        scale = self.action_scale.to(ml.device).expand(ml.shape[1] // 2).contiguous()
        bias = self.action_bias.to(ml.device).expand(ml.shape[1] // 2).contiguous()
        return _PolicyHead.apply(ml, eps, scale, bias)

    def to(self, device):
        self.action_scale = self.action_scale.to(device)
        self.action_bias = self.action_bias.to(device)
        return super().to(device)
