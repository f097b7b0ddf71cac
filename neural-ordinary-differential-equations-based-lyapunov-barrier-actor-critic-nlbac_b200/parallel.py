"""Data parallelism by trajectory: one process per GPU, weights replicated, rows sharded.

The path has exactly three cross-row reductions: parameter gradients, the scalar loss means, and the ``[m]``
constraint-row sums of ``get_policy_loss_2`` (U sac_cbf_clf.py:493-504), which must be combined BEFORE the
multiplier / ratio logic because that logic is non-linear in the sums (:507-528).  Convention used throughout:
every rank scales its local mean-type losses by ``rows_local / rows_global`` (``shard_weight``), combines constraint
sums with ``all_reduce_sum`` (identity backward), and gradients are all-reduced with SUM -- one flat buffer and ONE
collective per optimiser (``allreduce_grads``), after which every rank applies the identical optimiser step.

Transport: ``Comm`` wraps the C ABI's NCCL communicator (``nlbac_comm_*``, enqueued on the caller's CUDA stream);
without one the functions fall back to ``torch.distributed``'s default group (NCCL on GPUs, gloo in the CPU tests).
The reference's own analogue is the unused mpi4py scaffold ``utils/mpi_pytorch.py:20-34``.
"""
import ctypes

import torch

from . import _lib


def shard_rows(n_rows, world, rank):
    """Contiguous block of rows owned by ``rank``: the first ``n_rows % world`` ranks get one extra row."""
    base, extra = divmod(int(n_rows), int(world))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_weight(rows_local, rows_global):
    """Factor for a rank's local MEAN-type loss so that SUM-reduced gradients equal the global-batch gradient."""
    return float(rows_local) / float(rows_global)


class Comm:
    """NCCL communicator behind the C ABI.  ``Comm(world, rank)`` exchanges the 128-byte unique id through
    ``torch.distributed`` (any initialised backend, e.g. gloo) -- or pass ``unique_id`` bytes obtained elsewhere."""

    def __init__(self, world, rank, unique_id=None):
        L = _lib.lib()
        if unique_id is None:
            import torch.distributed as dist
            buf = ctypes.create_string_buffer(128)
            if rank == 0:
                _lib.check(L.nlbac_comm_unique_id(buf), "comm_unique_id")
            box = [bytes(buf.raw)]
            dist.broadcast_object_list(box, src=0)
            unique_id = box[0]
        self.world, self.rank = world, rank
        h = ctypes.c_void_p()
        _lib.check(L.nlbac_comm_init(world, rank, ctypes.c_char_p(unique_id), ctypes.byref(h)), "comm_init")
        self._h = h

    @staticmethod
    def make_unique_id():
        buf = ctypes.create_string_buffer(128)
        _lib.check(_lib.lib().nlbac_comm_unique_id(buf), "comm_unique_id")
        return bytes(buf.raw)

    def all_reduce_(self, flat):
        if not (flat.is_cuda and flat.dtype == torch.float32 and flat.is_contiguous()):
            raise RuntimeError("Comm.all_reduce_: contiguous float32 CUDA tensor expected")
        _lib.check(_lib.lib().nlbac_allreduce_grads(self._h, flat.data_ptr(), flat.numel(), _lib.raw_stream()),
                   "allreduce_grads")
        return flat

    def close(self):
        if self._h is not None:
            _lib.check(_lib.lib().nlbac_comm_destroy(self._h), "comm_destroy")
            self._h = None


def _all_reduce_(flat, comm):
    if comm is not None:
        return comm.all_reduce_(flat)
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(flat)
    return flat


def flat_view(grads):
    """One contiguous buffer holding ``grads``: the fused backward kernels already write every d-theta of a network
    into ONE allocation (the tensors are views of it), in which case nothing is copied."""
    base = grads[0]._base
    if base is not None and base.is_contiguous() and base.numel() == sum(g.numel() for g in grads) and all(
            g._base is base for g in grads):
        return base, False
    return torch.cat([g.reshape(-1) for g in grads]), True


@torch.no_grad()
def allreduce_grads(params, comm=None):
    """SUM all-reduce of ``p.grad`` over the ranks for every parameter that has one: one flat buffer, ONE collective
    (per call -- call it once per optimiser).  ``params``: a module, an optimiser or an iterable of parameters."""
    if isinstance(params, torch.nn.Module):
        params = list(params.parameters())
    elif isinstance(params, torch.optim.Optimizer):
        params = [p for gr in params.param_groups for p in gr["params"]]
    ps = [p for p in params if p.grad is not None]
    if not ps:
        return None
    grads = [p.grad for p in ps]
    flat, copied = flat_view(grads)
    _all_reduce_(flat, comm)
    if copied:
        off = 0
        for g in grads:
            g.copy_(flat[off:off + g.numel()].view_as(g))
            off += g.numel()
    return flat


def allreduce_grad_tensors(grads, comm=None):
    """Same for gradients returned by ``torch.autograd.grad``; returns the reduced flat buffer."""
    flat, copied = flat_view(list(grads))
    with torch.no_grad():
        _all_reduce_(flat, comm)
        if copied:
            off = 0
            for g in grads:
                g.copy_(flat[off:off + g.numel()].view_as(g))
                off += g.numel()
    return flat


class _AllReduceSum(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, comm):
        y = x.detach().clone().contiguous()
        _all_reduce_(y.view(-1), comm)
        return y

    @staticmethod
    def backward(ctx, g):
        # every rank evaluates the same L(sum_r x_r): dL/dx_r = dL/dsum, and SUM-reducing the parameter gradients
        # afterwards assembles dL/dsum * sum_r dx_r/dtheta without double counting
        return g, None


def all_reduce_sum(x, comm=None):
    """Differentiable SUM over the ranks, for the ``[m]`` constraint-row sums of a sharded minibatch."""
    return _AllReduceSum.apply(x, comm)


def train_step_dp(model, state, action, next_state, optimizer, loss_func, horizon, solver, time_interval, *,
                  rows_global, comm=None, time_batch=None, precision="fp32"):
    """``train_step`` (U model.py:221-260) on this rank's shard of a ``rows_global``-row NODE batch: fused solve + loss
    on the local rows, fused backward, ONE flat all-reduce of d-theta, the identical Adam step on every rank.
    Returns the GLOBAL loss / horizon like the reference (one host sync)."""
    from .odeint import odeint
    cols = (state, action) if time_batch is None else (state, action, time_batch)
    assert all(c.ndim == 2 for c in cols) and next_state.ndim == 2
    model.train()
    optimizer.zero_grad()
    pred = odeint(model, torch.cat(cols, dim=-1), [0.0, float(time_interval)], method=solver, atol=1e-7, rtol=1e-5,
                  precision=precision)[-1][:, :state.shape[1]]
    loss = loss_func(pred, next_state) * shard_weight(state.shape[0], rows_global)
    loss.backward()
    allreduce_grads(model, comm)
    optimizer.step()
    total = loss.detach().clone().reshape(1)
    _all_reduce_(total, comm)
    return total.item() / horizon
