"""Latency of the FP32 NODE solve alone at the reference's batch sizes (graph of 5 fwd+bwd, CUDA events)."""
import os, sys, statistics
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
import torch, nlbac_b200 as nb
def lat(fn, iters=100, warm=10):
    for _ in range(warm): fn()
    torch.cuda.synchronize(); ts = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1) * 1e3)
    return statistics.median(ts)
for tag, mk, D in (("U", lambda: nb.NeuralODEModel(3, 3, 6), 5), ("P", lambda: nb.NeuralODEModel(6, 6, 12), 8)):
    for B, method, steps in ((128, "euler", 1), (256, "euler", 1), (256, "rk4", 4)):
        torch.manual_seed(0)
        m = mk().cuda()
        y = torch.randn(B, D, device="cuda", requires_grad=True)
        ps = [y] + list(m.parameters())
        t = [0.02 * i for i in range(steps + 1)]
        def fb():
            return torch.autograd.grad(nb.odeint(m, y, t, method=method)[-1].sum(), ps)
        for _ in range(3): fb()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(5): keep = fb()
        print("%s B %4d %-5s x%d  fwd+bwd %.1f us" % (tag, B, method, steps, lat(g.replay) / 5))
