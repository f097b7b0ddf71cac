"""CUDA-graph capture of the NODE model-fit step.

At the reference's batch sizes one ``train_step`` (U model.py:221-260) is ~100 us of GPU work under ~500 us of
Python / autograd / launch overhead.  ``GraphedTrainStep`` captures the whole step once -- weight re-pack, fused
solve, loss, fused backward, Adam with its step counter on the device -- and replays it per call: one
``cudaGraphLaunch`` per model update.
"""
import ctypes

import torch

from . import _lib
from .odeint import odeint


class GraphedTrainStep:
    """``step = GraphedTrainStep(model, batch_size, n_s, n_u, solver=..., time_interval=...)`` then
    ``loss = step(state, action, next_state[, time_batch])`` -- the arithmetic of ``train_step`` with Adam
    (lr, default betas / eps), for a FIXED batch size.  Returns the loss as a 0-dim CUDA tensor (no host sync;
    the reference returns ``loss.item() / horizon``: divide and ``.item()`` it if that float is wanted).
    """

    def __init__(self, model, batch_size, n_s, n_u, *, solver="euler", time_interval=0.02, lr=1e-3,
                 betas=(0.9, 0.999), eps=1e-8, loss_func=None, has_time=False, precision="fp32", warmup=2):
        self.model = model
        self.params = [p for p in model.parameters()]
        dev = self.params[0].device
        if dev.type != "cuda":
            raise RuntimeError("GraphedTrainStep: the model must live on a CUDA device")
        self.lr, self.betas, self.eps = float(lr), betas, float(eps)
        self.loss_func = loss_func if loss_func is not None else torch.nn.functional.mse_loss
        self.n_s, self.solver, self.dt, self.precision = n_s, solver, float(time_interval), precision
        B = int(batch_size)
        self.state = torch.zeros(B, n_s, device=dev)
        self.action = torch.zeros(B, n_u, device=dev)
        self.next_state = torch.zeros(B, n_s, device=dev)
        self.time = torch.zeros(B, 1, device=dev) if has_time else None
        self.exp_avg = [torch.zeros_like(p) for p in self.params]
        self.exp_avg_sq = [torch.zeros_like(p) for p in self.params]
        self.step_dev = torch.zeros(2, dtype=torch.int32, device=dev)      # step count, ticket word of the Adam kernel
        if len(self.params) > _lib.MAX_TENSORS:
            raise RuntimeError("GraphedTrainStep: at most %d parameter tensors" % _lib.MAX_TENSORS)
        # eager warm-up on a side stream (lazy kernel attributes, allocator), then put everything back
        snap = [p.detach().clone() for p in self.params]
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(warmup):
                self._body()
        torch.cuda.current_stream().wait_stream(s)
        with torch.no_grad():
            for p, q in zip(self.params, snap):
                p.copy_(q)
            for t in self.exp_avg + self.exp_avg_sq:
                t.zero_()
            self.step_dev.zero_()
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.loss = self._body()

    def _body(self):
        cols = (self.state, self.action) if self.time is None else (self.state, self.action, self.time)
        y0 = torch.cat(cols, dim=-1)
        pred = odeint(self.model, y0, [0.0, self.dt], method=self.solver, atol=1e-7, rtol=1e-5,
                      precision=self.precision)[-1][:, :self.n_s]
        loss = self.loss_func(pred, self.next_state)
        grads = [g.contiguous() for g in torch.autograd.grad(loss, self.params)]
        n = len(self.params)
        arr = lambda ts: (ctypes.c_void_p * n)(*[t.data_ptr() for t in ts])  # noqa: E731
        _lib.check(_lib.lib().nlbac_adam_step_multi_dev(
            n, arr(self.params), arr(grads), arr(self.exp_avg), arr(self.exp_avg_sq),
            (ctypes.c_int64 * n)(*[p.numel() for p in self.params]), self.step_dev.data_ptr(), 1, self.lr,
            self.betas[0], self.betas[1], self.eps, _lib.raw_stream()), "adam_step_multi_dev")
        for p in self.params:
            torch.autograd.graph.increment_version(p)
        return loss

    @torch.no_grad()
    def __call__(self, state, action, next_state, time_batch=None):
        if state.shape != self.state.shape:
            raise ValueError("GraphedTrainStep was captured for batch %s, got %s" % (tuple(self.state.shape),
                                                                                     tuple(state.shape)))
        self.state.copy_(state, non_blocking=True)
        self.action.copy_(action, non_blocking=True)
        self.next_state.copy_(next_state, non_blocking=True)
        if self.time is not None:
            self.time.copy_(time_batch, non_blocking=True)
        self.graph.replay()
        for p in self.params:       # the replay wrote the weights through raw pointers: invalidate packed caches
            torch.autograd.graph.increment_version(p)
        return self.loss
