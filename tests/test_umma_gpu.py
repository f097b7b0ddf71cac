"""tcgen05 kind::tf32 operand-layout self-test (K-major / MN-major interpretations of the same shared-memory
images) against a CPU reference with operands rounded to TF32.  Tolerance 2e-3 relative (TF32 truncation vs RN)."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_umma_layouts(mode):
    import nlbac_b200
    L = nlbac_b200._lib.lib()
    g = torch.Generator().manual_seed(mode)
    X = torch.randn(128, 104, generator=g)
    Y = torch.randn(128 if mode == 2 else 104, 104, generator=g)
    D = torch.full((128, 112), float("nan"), device="cuda")
    Xc, Yc = X.cuda(), Y.cuda()
    nlbac_b200._lib.check(L.nlbac_umma_selftest(mode, Xc.data_ptr(), Yc.data_ptr(), D.data_ptr(),
                                                torch.cuda.current_stream().cuda_stream), "umma_selftest")
    torch.cuda.synchronize()
    Xd, Yd = X.double(), Y.double()
    want = Xd @ Yd.t() if mode == 0 else (Xd @ Yd if mode == 1 else Xd.t() @ Yd)
    got = D.cpu().double()[:want.shape[0], :want.shape[1]]
    err = ((got - want).abs().max() / want.abs().max()).item()
    assert err < 2e-3, err
