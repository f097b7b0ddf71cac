"""Adaptive dopri5 behind ``odeint(..., method='dopri5')`` (torchdiffeq's default method).

torchdiffeq semantics kept: batch-global RMS error norm -> ONE step size for the
whole batch, ``_select_initial_step``, ``_optimal_step_size`` (safety 0.9, ifactor 10, dfactor 0.2), FSAL,
no clipping of the last step, 4th-order dense output.  The step sizes are constants of the backward pass.

Two controllers with the identical step sequence (same double arithmetic):
  * ``device`` (default): accept / reject and the next step size are decided by a one-thread kernel after each
    attempt; the host reads the state back ONCE per batch of enqueued attempts (eager) or never (under CUDA-graph
    capture, where ``options["max_attempts"]`` attempts are recorded and the surplus ones are no-op launches).
    Sharded batch (one process per GPU): ``options={"comm": parallel.Comm, "rows_global": N}`` all-reduces the three
    squared-norm sums before every controller step, so every rank takes the step sequence of the unsharded batch.
  * ``host`` (``options={"controller": "host"}``): one scalar read-back per attempted step, exactly where
    torchdiffeq itself synchronises (rk_common._adaptive_step).
"""
import ctypes

import numpy as np
import torch

from . import _lib

DEFAULT_STEP_CAPACITY = 128
FIRST_ATTEMPT_BATCH = 8          # eager device controller: attempts enqueued before the first read-back
DEFAULT_CAPTURE_ATTEMPTS = 16    # under graph capture nothing can be read back: a fixed attempt budget
_ATTEMPT_HINT = {}               # (B, D, t, rtol, atol) -> attempts the last eager solve of that problem needed
ERRORS = {1: "non-finite error ratio in dopri5", 2: "dopri5 exceeded max_steps accepted steps",
          3: "dopri5 stalled (rejected steps)", 4: "underflow in dt",
          5: "dopri5 did not finish within the enqueued attempts"}


def _grads_from_flat(gflat, shapes):
    grads, off = [], 0
    for shp in shapes:
        n = int(np.prod(shp))
        grads.append(gflat[off:off + n].view(shp))
        off += n
    return grads


class DeviceStats(dict):
    """Statistics of a device-driven solve that was never read back (graph capture): ``read()`` synchronises."""

    def read(self):
        L = _lib.lib()
        st, n_t, cap = self["_state"], self["_n_t"], self["_cap"]
        na, nr, dn, er = ctypes.c_int(), ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        dts = np.zeros(cap, dtype=np.float32)
        _lib.check(L.nlbac_node_dopri5_state_read(st.data_ptr(), n_t, cap, ctypes.byref(na), ctypes.byref(nr),
                                                  ctypes.byref(dn), ctypes.byref(er),
                                                  dts.ctypes.data_as(ctypes.POINTER(ctypes.c_float)),
                                                  _lib.raw_stream()), "node_dopri5_state_read")
        self.update(n_accepted=na.value, n_rejected=nr.value, done=dn.value, err=er.value,
                    nfe=2 + 6 * (na.value + nr.value), step_dt=dts[:na.value].copy())
        return self


class _Dopri5DeviceSolve(torch.autograd.Function):
    @staticmethod
    def forward(ctx, y0, desc, packed, t_np, rtol, atol, precision, want_ckpt, max_steps, stats, attempts, shard,
                *params):
        L = _lib.lib()
        B, D = y0.shape
        n_t = len(t_np)
        dev = y0.device
        y0c = y0.detach().contiguous()
        fwd, ck, bwd, sb = ctypes.c_size_t(), ctypes.c_size_t(), ctypes.c_size_t(), ctypes.c_size_t()
        _lib.check(L.nlbac_node_dopri5_workspace_bytes(ctypes.byref(desc), B, n_t, max_steps, ctypes.byref(fwd),
                                                       ctypes.byref(ck), ctypes.byref(bwd)), "dopri5_workspace_bytes")
        _lib.check(L.nlbac_node_dopri5_state_bytes(n_t, max_steps, ctypes.byref(sb)), "dopri5_state_bytes")
        ws = torch.empty(fwd.value, dtype=torch.uint8, device=dev)
        state = torch.empty((sb.value + 7) // 8, dtype=torch.int64, device=dev)
        ckpt = torch.empty(max(ck.value // 4, 1), dtype=torch.float32, device=dev) if want_ckpt else None
        y_out = torch.empty((n_t, B, D), dtype=torch.float32, device=dev)
        t_arr = np.ascontiguousarray(t_np, dtype=np.float32)
        fp = ctypes.POINTER(ctypes.c_float)
        capturing = torch.cuda.is_current_stream_capturing()
        comm_h, rows_global = (shard[0]._h, int(shard[1])) if shard is not None else (None, B)

        def enqueue(n_attempts, resume):
            _lib.check(L.nlbac_node_dopri5_fwd_dev(
                ctypes.byref(desc), packed.data_ptr(), y0c.data_ptr(), B, t_arr.ctypes.data_as(fp), n_t, rtol, atol,
                max_steps, n_attempts, resume, precision, y_out.data_ptr(), ckpt.data_ptr() if want_ckpt else None,
                ws.data_ptr(), state.data_ptr(), comm_h, rows_global, _lib.raw_stream()), "node_dopri5_fwd_dev")

        info = DeviceStats(_state=state, _n_t=n_t, _cap=max_steps)
        hint_key = (rows_global, D, tuple(float(v) for v in t_arr), rtol, atol)
        if capturing:
            # nothing can be read back while recording: a fixed budget -- the caller's, else twice what the eager
            # warm-up calls of the same problem needed (+2), else the default.  Surplus attempts are no-op launches.
            hint = _ATTEMPT_HINT.get(hint_key)
            enqueue(attempts if attempts > 0 else (2 * hint + 2 if hint else DEFAULT_CAPTURE_ATTEMPTS), 0)
        else:
            hint = _ATTEMPT_HINT.get(hint_key)
            batch = attempts if attempts > 0 else (hint + 1 if hint else FIRST_ATTEMPT_BATCH)
            enqueue(batch, 0)
            while True:
                info.read()
                if info["done"]:
                    break
                batch = min(2 * batch, 64)
                enqueue(batch, 1)
            _ATTEMPT_HINT[hint_key] = info["n_accepted"] + info["n_rejected"]
            if info["err"]:
                raise RuntimeError(f"{ERRORS.get(info['err'], 'dopri5 error')} (max_num_steps={max_steps}, "
                                   f"accepted {info['n_accepted']}, rejected {info['n_rejected']})")
        if stats is not None:
            stats.update(info)

            def read():          # after a captured solve has been replayed: synchronise and refresh ``stats``
                stats.update(info.read())
                return stats
            stats["read"] = read
        if want_ckpt:
            ctx.save_for_backward(y0c, packed, ckpt, state)
            ctx.meta = (desc, n_t, max_steps, precision, bwd.value, [p.shape for p in params])
        return y_out

    @staticmethod
    def backward(ctx, grad_out):
        L = _lib.lib()
        y0c, packed, ckpt, state = ctx.saved_tensors
        desc, n_t, max_steps, precision, bwd_bytes, shapes = ctx.meta
        B, D = y0c.shape
        dev = y0c.device
        need_y0 = ctx.needs_input_grad[0]
        need_p = any(ctx.needs_input_grad[12:])
        go = grad_out.contiguous()
        ws = torch.empty(max(bwd_bytes, 16), dtype=torch.uint8, device=dev)
        gy0 = torch.empty((B, D), dtype=torch.float32, device=dev) if need_y0 else None
        n_param = L.nlbac_node_param_count(ctypes.byref(desc))
        gflat = torch.empty(n_param, dtype=torch.float32, device=dev) if need_p else None
        _lib.check(L.nlbac_node_dopri5_bwd_dev(
            ctypes.byref(desc), packed.data_ptr(), y0c.data_ptr(), B, n_t, max_steps, state.data_ptr(), precision,
            ckpt.data_ptr(), go.data_ptr(), gy0.data_ptr() if need_y0 else None,
            gflat.data_ptr() if need_p else None, ws.data_ptr(), _lib.raw_stream()), "node_dopri5_bwd_dev")
        grads = _grads_from_flat(gflat, shapes) if need_p else [None] * len(shapes)
        return (gy0, None, None, None, None, None, None, None, None, None, None, None, *grads)


class _Dopri5Solve(torch.autograd.Function):
    @staticmethod
    def forward(ctx, y0, desc, packed, t_np, rtol, atol, precision, want_ckpt, max_steps, stats, *params):
        L = _lib.lib()
        B, D = y0.shape
        n_t = len(t_np)
        dev = y0.device
        y0c = y0.detach().contiguous()
        fwd, ck, bwd = ctypes.c_size_t(), ctypes.c_size_t(), ctypes.c_size_t()
        _lib.check(L.nlbac_node_dopri5_workspace_bytes(ctypes.byref(desc), B, n_t, max_steps, ctypes.byref(fwd),
                                                       ctypes.byref(ck), ctypes.byref(bwd)), "dopri5_workspace_bytes")
        ws = torch.empty(fwd.value, dtype=torch.uint8, device=dev)
        ckpt = torch.empty(max(ck.value // 4, 1), dtype=torch.float32, device=dev) if want_ckpt else None
        y_out = torch.empty((n_t, B, D), dtype=torch.float32, device=dev)
        t_arr = np.ascontiguousarray(t_np, dtype=np.float32)
        step_dt = np.zeros(max_steps, dtype=np.float32)
        out_step = np.zeros(n_t, dtype=np.int32)
        out_x = np.zeros(n_t, dtype=np.float32)
        n_acc, n_rej = ctypes.c_int(), ctypes.c_int()
        fp, ip = ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_int)
        _lib.check(L.nlbac_node_dopri5_fwd(
            ctypes.byref(desc), packed.data_ptr(), y0c.data_ptr(), B, t_arr.ctypes.data_as(fp), n_t, rtol, atol,
            max_steps, precision, y_out.data_ptr(), ckpt.data_ptr() if want_ckpt else None, ws.data_ptr(),
            ctypes.byref(n_acc), ctypes.byref(n_rej), step_dt.ctypes.data_as(fp), out_step.ctypes.data_as(ip),
            out_x.ctypes.data_as(fp), _lib.raw_stream()), "node_dopri5_fwd")
        if stats is not None:
            stats.update(n_accepted=n_acc.value, n_rejected=n_rej.value, nfe=2 + 6 * (n_acc.value + n_rej.value),
                         step_dt=step_dt[:n_acc.value].copy())
        if want_ckpt:
            ctx.save_for_backward(y0c, packed, ckpt)
            ctx.meta = (desc, n_t, n_acc.value, step_dt, out_step, out_x, precision, bwd.value,
                        [p.shape for p in params])
        return y_out

    @staticmethod
    def backward(ctx, grad_out):
        L = _lib.lib()
        y0c, packed, ckpt = ctx.saved_tensors
        desc, n_t, n_acc, step_dt, out_step, out_x, precision, bwd_bytes, shapes = ctx.meta
        B, D = y0c.shape
        dev = y0c.device
        need_y0 = ctx.needs_input_grad[0]
        need_p = any(ctx.needs_input_grad[10:])
        go = grad_out.contiguous()
        ws = torch.empty(max(bwd_bytes, 16), dtype=torch.uint8, device=dev)
        gy0 = torch.empty((B, D), dtype=torch.float32, device=dev) if need_y0 else None
        n_param = L.nlbac_node_param_count(ctypes.byref(desc))
        gflat = torch.empty(n_param, dtype=torch.float32, device=dev) if need_p else None
        fp, ip = ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_int)
        _lib.check(L.nlbac_node_dopri5_bwd(
            ctypes.byref(desc), packed.data_ptr(), y0c.data_ptr(), B, n_t, n_acc, step_dt.ctypes.data_as(fp),
            out_step.ctypes.data_as(ip), out_x.ctypes.data_as(fp), precision, ckpt.data_ptr(), go.data_ptr(),
            gy0.data_ptr() if need_y0 else None, gflat.data_ptr() if need_p else None, ws.data_ptr(),
            _lib.raw_stream()), "node_dopri5_bwd")
        grads = [None] * len(shapes)
        if need_p:
            off = 0
            for i, shp in enumerate(shapes):
                n = int(np.prod(shp))
                grads[i] = gflat[off:off + n].view(shp)
                off += n
        return (gy0, None, None, None, None, None, None, None, None, None, *grads)


def dopri5_solve(func, y0, desc, packed, params, t_np, rtol, atol, precision, need_grad, options):
    cap = int(options.pop("max_num_steps", DEFAULT_STEP_CAPACITY))
    stats = options.pop("_stats", None)
    controller = options.pop("controller", "device")
    attempts = int(options.pop("max_attempts", 0))
    if controller not in ("device", "host"):
        raise ValueError(f"dopri5 controller {controller!r}: 'device' or 'host'")
    comm, rows_global = options.pop("comm", None), options.pop("rows_global", None)
    if (comm is None) != (rows_global is None):
        raise ValueError("dopri5 on a sharded batch needs both options: comm (parallel.Comm) and rows_global")
    if controller == "device":
        shard = (comm, rows_global) if comm is not None else None
        return _Dopri5DeviceSolve.apply(y0, desc, packed, t_np, rtol, atol, precision, need_grad, cap, stats, attempts,
                                        shard, *params)
    if comm is not None:
        raise ValueError("dopri5 on a sharded batch runs with the device-side controller only")
    return _Dopri5Solve.apply(y0, desc, packed, t_np, rtol, atol, precision, need_grad, cap, stats, *params)
