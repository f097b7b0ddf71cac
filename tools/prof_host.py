"""Where the host time of one eager odeint forward + backward at B = 128 goes (cProfile, 300 calls)."""
import os, sys, cProfile, pstats, time
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
import torch, nlbac_b200 as nb
torch.manual_seed(0)
m = nb.NeuralODEModel(3, 3, 6).cuda()
y = torch.randn(128, 5, device="cuda")
ps = list(m.parameters())
def fb():
    sol = nb.odeint(m, y, [0.0, 0.02], method="euler", atol=1e-7, rtol=1e-5)
    return torch.autograd.grad(sol[-1][:, :3].sum(), ps)
for _ in range(20): fb()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(300): fb()
torch.cuda.synchronize()
print("us per fwd+bwd (wall, no profiler): %.1f" % ((time.perf_counter() - t0) / 300 * 1e6))
pr = cProfile.Profile(); pr.enable()
for _ in range(300): fb()
torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(22)
