"""Latency of the fused MLP forward / backward alone (CUDA events, median of 200) at the reference's batch sizes."""
import os, sys, statistics
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
import torch, nlbac_b200 as nb
def lat(fn, iters=200, warm=20):
    for _ in range(warm): fn()
    torch.cuda.synchronize(); ts = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1) * 1e3)
    return statistics.median(ts)
for B in (128, 256, 4096, 65536):
    net = nb.LyaNetwork(11, 256).cuda()
    x = torch.randn(B, 11, device="cuda")
    g = torch.cuda.CUDAGraph()
    with torch.no_grad():
        net(x); torch.cuda.synchronize()
        with torch.cuda.graph(g):
            for _ in range(10): y = net(x)
    print("B %6d  fwd (graph of 10) %.2f us each" % (B, lat(g.replay) / 10))
for B in (128, 256, 4096, 65536):
    net = nb.LyaNetwork(11, 256).cuda()
    x = torch.randn(B, 11, device="cuda", requires_grad=True)
    ps = [x] + list(net.parameters())
    def fb():
        return torch.autograd.grad(net(x).sum(), ps)
    for _ in range(3): fb()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(5): keep = fb()
    print("B %6d  fwd+bwd (graph of 5) %.2f us each" % (B, lat(g.replay) / 5))
