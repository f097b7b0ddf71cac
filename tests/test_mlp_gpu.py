"""GPU parity of the fused actor / critic / Lyapunov / barrier MLPs (nlbac_mlp_*, nlbac_policy_head_*) against
the golden vectors produced by the reference's own classes and against the CPU oracle.  Tolerance rel 1e-5
(fp32 FMA arithmetic, different summation order than cuBLAS), max-norm per tensor."""
import os

import numpy as np
import pytest
import torch

from oracle import nlbac_ref

pytestmark = pytest.mark.gpu
TOL = 1e-5


def _rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


class _Box:
    def __init__(self, hi):
        self.high = np.array(hi, dtype="float32")
        self.low = -self.high


@pytest.mark.parametrize("tag", ["U", "C", "BP"])
def test_mlps_match_reference_golden(tag, golden_dir):
    import nlbac_b200
    g = torch.load(os.path.join(golden_dir, "mlp_%s.pt" % tag), weights_only=False)
    obs, act, H = g["obs"].cuda(), g["act"].cuda(), g["hidden"]
    n_o, n_u = obs.shape[1], act.shape[1]
    # twin Q
    q = nlbac_b200.QNetwork(n_o, n_u, H)
    q.load_state_dict(g["q"]["state_dict"])
    q.cuda()
    og, ag = obs.clone().requires_grad_(), act.clone().requires_grad_()
    q1, q2 = q(og, ag)
    assert _rel(q1, g["q"]["q1"]) <= TOL and _rel(q2, g["q"]["q2"]) <= TOL
    ((q1 * g["q"]["c1"].cuda()).sum() + (q2 * g["q"]["c2"].cuda()).sum()).backward()
    assert _rel(og.grad, g["q"]["grad_obs"]) <= TOL and _rel(ag.grad, g["q"]["grad_act"]) <= TOL
    for k, p in q.named_parameters():
        assert _rel(p.grad, g["q"]["grad_params"][k]) <= TOL, k
    # Lyapunov
    lya = nlbac_b200.LyaNetwork(g["lya_in"].shape[1], H)
    lya.load_state_dict(g["lya"]["state_dict"])
    lya.cuda()
    li = g["lya_in"].cuda().requires_grad_()
    v = lya(li)
    assert _rel(v, g["lya"]["v"]) <= TOL
    (v * g["lya"]["c"].cuda()).sum().backward()
    assert _rel(li.grad, g["lya"]["grad_in"]) <= TOL
    for k, p in lya.named_parameters():
        assert _rel(p.grad, g["lya"]["grad_params"][k]) <= TOL, k
    # barrier
    if "barrier" in g:
        bar = nlbac_b200.BarrierNetwork(n_o, n_u, H)
        bar.load_state_dict(g["barrier"]["state_dict"])
        bar.cuda()
        og, ag = obs.clone().requires_grad_(), act.clone().requires_grad_()
        b = bar(og, ag)
        assert _rel(b, g["barrier"]["b"]) <= TOL
        (b * g["barrier"]["c"].cuda()).sum().backward()
        assert _rel(og.grad, g["barrier"]["grad_obs"]) <= TOL and _rel(ag.grad, g["barrier"]["grad_act"]) <= TOL
        for k, p in bar.named_parameters():
            assert _rel(p.grad, g["barrier"]["grad_params"][k]) <= TOL, k
    # policy (tanh-Gaussian head, injected eps)
    pol = nlbac_b200.GaussianPolicy(n_o, n_u, H, _Box(g["action_high"]))
    pol.load_state_dict(g["policy"]["state_dict"])
    pol.to(torch.device("cuda"))
    og = obs.clone().requires_grad_()
    a, logp, mean = pol.sample(og, eps=g["eps"].cuda())
    assert _rel(a, g["policy"]["action"]) <= TOL and _rel(mean, g["policy"]["mean"]) <= TOL
    assert _rel(logp, g["policy"]["log_prob"]) <= 2 * TOL
    ((a * g["policy"]["ca"].cuda()).sum() + (logp * g["policy"]["cl"].cuda()).sum()).backward()
    assert _rel(og.grad, g["policy"]["grad_obs"]) <= 2 * TOL
    for k, p in pol.named_parameters():
        assert _rel(p.grad, g["policy"]["grad_params"][k]) <= 2 * TOL, k
    m2, ls2 = pol(obs)
    assert m2.shape == (obs.shape[0], n_u) and ls2.min() >= -20 and ls2.max() <= 2


@pytest.mark.parametrize("B", [1, 7, 256, 1000, 4095, 4096, 70001])
def test_mlp_ragged_batches_vs_oracle(B):
    """Row tiles of 8: batches that are not a multiple of the tile, B=1, and more rows than one wave; B >= 4096 takes
    the split-over-the-batch parameter-gradient kernels (mlp_atb_split_kernel / mlp_colsum_split_kernel), with ragged
    chunk ends."""
    import nlbac_b200
    torch.manual_seed(B)
    ref = nlbac_ref.LyaNetwork(11, 256)
    with torch.no_grad():
        for p in ref.parameters():
            if p.ndim == 1:
                p.normal_(0, 0.1)
    ours = nlbac_b200.LyaNetwork(11, 256)
    ours.load_state_dict(ref.state_dict())
    ours.cuda()
    x = torch.randn(B, 11)
    cot = torch.randn(B, 1)
    xr = x.clone().requires_grad_()
    (ref(xr) * cot).sum().backward()
    xo = x.cuda().requires_grad_()
    out = ours(xo)
    assert _rel(out, ref(x)) <= TOL
    (out * cot.cuda()).sum().backward()
    if B < 4096:
        assert _rel(xo.grad, xr.grad) <= TOL
    else:   # kink rows (test_node_gpu._check): a pre-activation within fp32 round-off of zero flips that ROW's mask
        d = (xo.grad.detach().cpu().double() - xr.grad.double()).abs().max(1).values
        bad = int((d > TOL * xr.grad.abs().max().item()).sum())
        assert bad <= max(2, B // 10000), ("kink rows", bad)
    # fp64 tie-breaker (as in test_node_gpu._check): with tens of thousands of rows and a random cotangent the parameter
    # gradients are cancelling sums, and the fp32 CPU oracle itself sits 1e-4 .. 1e-3 from fp64 (tools/diag_mlp_bigB.py:
    # ours and the fp32 oracle are equally far from fp64) -- the kernel may not be further from fp64 than 2x the oracle is
    r64 = nlbac_ref.LyaNetwork(11, 256).double()
    r64.load_state_dict({k: v.double() for k, v in ref.state_dict().items()})
    (r64(x.double()) * cot.double()).sum().backward()
    for (k, p), (_, r), (_, d) in zip(ours.named_parameters(), ref.named_parameters(), r64.named_parameters()):
        e32 = _rel(p.grad, r.grad)
        assert e32 <= TOL or _rel(p.grad, d.grad) <= max(TOL, 2 * _rel(r.grad, d.grad)), (k, e32)


def test_mlp_empty_batch_and_errors():
    import nlbac_b200
    net = nlbac_b200.LyaNetwork(4, 256).cuda()
    out = net(torch.zeros(0, 4, device="cuda"))
    assert out.shape == (0, 1)
    with pytest.raises(RuntimeError):
        nlbac_b200.LyaNetwork(4, 256)(torch.zeros(3, 4))          # CPU tensor: no fallback
    with pytest.raises(RuntimeError):
        nlbac_b200.LyaNetwork(4, 100).cuda()(torch.zeros(3, 4, device="cuda"))   # hidden not a multiple of 32


def test_policy_log_std_clamp_blocks_gradient():
    """Rows whose raw log_std falls outside [-20, 2] get no gradient through log_std (torch.clamp semantics)."""
    from nlbac_b200.mlp import _PolicyHead
    ml = torch.tensor([[0.3, 5.0], [0.1, -25.0], [0.2, 0.5]], device="cuda", requires_grad=True)
    eps = torch.tensor([[0.5], [0.5], [0.5]], device="cuda")
    one, zero = torch.ones(1, device="cuda"), torch.zeros(1, device="cuda")
    a, lp, _ = _PolicyHead.apply(ml, eps, one, zero)
    (a.sum() + lp.sum()).backward()
    assert ml.grad[0, 1] == 0 and ml.grad[1, 1] == 0 and ml.grad[2, 1] != 0
    m = ml.detach().cpu().requires_grad_()
    ls = torch.clamp(m[:, 1:], -20, 2)
    x = m[:, :1] + ls.exp() * eps.cpu()
    y = torch.tanh(x)
    lp_ref = (-((x - m[:, :1]) ** 2) / (2 * ls.exp() ** 2) - ls - 0.5 * np.log(2 * np.pi) - torch.log(1 - y ** 2 + 1e-6))
    (y.sum() + lp_ref.sum()).backward()
    assert _rel(ml.grad, m.grad) <= 5e-5   # saturated tanh: 1 - y^2 cancels, tanhf is ~2 ulp


@pytest.mark.parametrize("H", [64, 128, 512])
def test_mlp_other_hidden_sizes(H):
    """--hidden_size other than the default 256 (U main.py:216): any multiple of 32 in [32, 1024] is accepted."""
    import nlbac_b200
    torch.manual_seed(H)
    ref = nlbac_ref.QNetwork(11, 2, H)
    ours = nlbac_b200.QNetwork(11, 2, H)
    ours.load_state_dict(ref.state_dict())
    ours.cuda()
    obs, act = torch.randn(200, 11), torch.randn(200, 2)
    ar = act.clone().requires_grad_()
    r1, r2 = ref(obs, ar)
    (r1.sum() + 2 * r2.sum()).backward()
    ao = act.cuda().requires_grad_()
    q1, q2 = ours(obs.cuda(), ao)
    assert _rel(q1, r1) <= TOL and _rel(q2, r2) <= TOL
    (q1.sum() + 2 * q2.sum()).backward()
    assert _rel(ao.grad, ar.grad) <= TOL
    for (k, p), (_, r) in zip(ours.named_parameters(), ref.named_parameters()):
        assert _rel(p.grad, r.grad) <= TOL, k
    with pytest.raises(RuntimeError):
        nlbac_b200.QNetwork(11, 2, 100).cuda()(obs.cuda(), act.cuda())      # 100 is not a multiple of 32
