"""nlbac_node_solve_bwd_noalloc: the FP32 reverse sweep with its scratch in the library's static device array
(include/nlbac.h; no counterpart in the reference).

Collected last on purpose.  The new entry point is called through the C ABI and held
to the EXISTING entry point (nlbac_node_solve_bwd, caller-supplied workspace) on the same inputs, on two calls in a row
with different batches, so the second call runs over the first call's leftovers in the shared array.  Tolerance: the
fp32 bar of tests/test_node_gpu.py (rel 1e-5 per tensor; the two calls reduce the same terms with atomics in a
different order, they are not bit-identical).
"""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
TOL = 1e-5


def _rel(a, b):
    a, b = a.double(), b.double()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def _solve_both(L, lib, desc, packed, n_param, y0, t_np, cot):
    B, D = y0.shape
    n_t = len(t_np)
    tp = t_np.ctypes.data_as(ctypes.POINTER(ctypes.c_float))
    st = torch.cuda.current_stream().cuda_stream
    fwd, ck, bwd = ctypes.c_size_t(), ctypes.c_size_t(), ctypes.c_size_t()
    lib.check(L.nlbac_node_workspace_bytes(ctypes.byref(desc), B, n_t, n_t, 2, ctypes.byref(fwd), ctypes.byref(ck),
                                           ctypes.byref(bwd)), "workspace_bytes")
    ws_f = torch.empty(max(fwd.value, 16), dtype=torch.uint8, device="cuda")
    ckpt = torch.empty(ck.value // 4, dtype=torch.float32, device="cuda")
    y_out = torch.empty((n_t, B, D), dtype=torch.float32, device="cuda")
    lib.check(L.nlbac_node_solve_fwd(ctypes.byref(desc), packed.data_ptr(), y0.data_ptr(), B, tp, n_t, tp, n_t, 2, 0,
                                     y_out.data_ptr(), ckpt.data_ptr(), ws_f.data_ptr(), st), "solve_fwd")
    # existing entry point: caller-supplied workspace
    ws_b = torch.empty(bwd.value, dtype=torch.uint8, device="cuda")
    gy_a = torch.empty((B, D), dtype=torch.float32, device="cuda")
    gp_a = torch.empty(n_param, dtype=torch.float32, device="cuda")
    lib.check(L.nlbac_node_solve_bwd(ctypes.byref(desc), packed.data_ptr(), y0.data_ptr(), B, tp, n_t, tp, n_t, 2, 0,
                                     ckpt.data_ptr(), cot.data_ptr(), gy_a.data_ptr(), gp_a.data_ptr(),
                                     ws_b.data_ptr(), st), "solve_bwd")
    # new entry point: outputs allocated by the caller BEFORE the call; the call itself must not allocate
    gy_b = torch.full((B, D), float("nan"), dtype=torch.float32, device="cuda")
    gp_b = torch.full((n_param,), float("nan"), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    lib.check(L.nlbac_node_solve_bwd_noalloc(ctypes.byref(desc), packed.data_ptr(), y0.data_ptr(), B, tp, n_t, tp, n_t,
                                             2, ckpt.data_ptr(), cot.data_ptr(), gy_b.data_ptr(), gp_b.data_ptr(), st),
              "solve_bwd_noalloc")
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    return (gy_a, gp_a), (gy_b, gp_b)


def test_noalloc_backward_matches_existing_entry_point_twice():
    import nlbac_b200
    from nlbac_b200 import _lib as lib
    from nlbac_b200.odeint import describe, _packed
    L = lib.lib()
    torch.manual_seed(0)
    model = nlbac_b200.NeuralODEModel(3, 3, 6).cuda()
    desc, params = describe(model, 5)
    packed = _packed(model, desc, params)
    n_param = L.nlbac_node_param_count(ctypes.byref(desc))
    t_np = np.ascontiguousarray(0.02 * np.arange(4), dtype=np.float32)
    g = torch.Generator(device="cuda").manual_seed(3)
    # two calls in a row: different batch sizes and cotangents, the second over the first one's leftovers
    for B in (777, 130):
        y0 = torch.randn(B, 5, generator=g, device="cuda")
        cot = torch.randn(len(t_np), B, 5, generator=g, device="cuda")
        (gy_a, gp_a), (gy_b, gp_b) = _solve_both(L, lib, desc, packed, n_param, y0, t_np, cot)
        assert torch.isfinite(gy_b).all() and torch.isfinite(gp_b).all()
        assert _rel(gy_b, gy_a) <= TOL, (B, _rel(gy_b, gy_a))
        assert _rel(gp_b, gp_a) <= TOL, (B, _rel(gp_b, gp_a))


def test_noalloc_rejects_oversized_and_reports():
    import nlbac_b200
    from nlbac_b200 import _lib as lib
    from nlbac_b200.odeint import describe, _packed
    L = lib.lib()
    model = nlbac_b200.NeuralODEModel(3, 3, 6).cuda()
    desc, params = describe(model, 5)
    packed = _packed(model, desc, params)
    n_t = 200000                      # schedule alone (12 bytes per output time) exceeds the 2 MiB array
    t_np = np.ascontiguousarray(1e-4 * np.arange(n_t), dtype=np.float32)
    tp = t_np.ctypes.data_as(ctypes.POINTER(ctypes.c_float))
    dummy = torch.zeros(16, device="cuda")
    rc = L.nlbac_node_solve_bwd_noalloc(ctypes.byref(desc), packed.data_ptr(), dummy.data_ptr(), 1, tp, n_t, tp, n_t, 2,
                                        dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(), None,
                                        torch.cuda.current_stream().cuda_stream)
    assert rc == -1
    msg = L.nlbac_last_error().decode()
    assert "scratch" in msg and "nlbac_node_solve_bwd" in msg, msg
