"""tcgen05 operand-layout self-test: kind::tf32 with K-major operands (hidden-layer and dX products) and
kind::f16 with MN-major bf16 images (dW products) against a CPU product of the same rounded operands.
Tolerance 2e-3 relative for TF32 (hardware truncation vs round-to-nearest), 1e-5 for bf16 (operands pre-rounded,
fp32 accumulation)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_umma_layouts(mode):
    import nlbac_b200
    L = nlbac_b200._lib.lib()
    g = torch.Generator().manual_seed(mode)
    X = torch.randn(128, 104, generator=g)
    Y = torch.randn(128 if mode == 1 else 104, 104, generator=g)
    D = torch.full((128, 112), float("nan"), device="cuda")
    Xc, Yc = X.cuda(), Y.cuda()
    nlbac_b200._lib.check(L.nlbac_umma_selftest(mode, Xc.data_ptr(), Yc.data_ptr(), D.data_ptr(),
                                                torch.cuda.current_stream().cuda_stream), "umma_selftest")
    torch.cuda.synchronize()
    if mode in (0, 2):
        want, tol = X.double() @ Y.double().t(), 2e-3
    else:
        want, tol = X.bfloat16().double().t() @ Y.bfloat16().double(), 1e-5
    got = D.cpu().double()[:want.shape[0], :want.shape[1]]
    err = ((got - want).abs().max() / want.abs().max()).item()
    assert err < tol, err
