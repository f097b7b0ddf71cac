"""cProfile of the host-side path for a launch-bound call (B = 128 Unicycle euler forward + backward)."""
import cProfile, pstats, sys, os, io
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import nlbac_b200
m = nlbac_b200.NeuralODEModel(3, 3, 6).cuda()
y = torch.rand(128, 5, device="cuda")
ps = list(m.parameters())
def one():
    sol = nlbac_b200.odeint(m, y, [0.0, 0.02], method="euler", atol=1e-7, rtol=1e-5)
    torch.autograd.grad(sol[-1][:, :3].sum(), ps)
for _ in range(20): one()
torch.cuda.synchronize()
pr = cProfile.Profile(); pr.enable()
for _ in range(300): one()
torch.cuda.synchronize(); pr.disable()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(45); print(s.getvalue()[:9000])
q = nlbac_b200.QNetwork(7, 2, 256).cuda(); xs = (torch.randn(128, 7, device="cuda"), torch.randn(128, 2, device="cuda")); qp = list(q.parameters())
def two():
    a, b = q(*xs); torch.autograd.grad(a.sum() + b.sum(), qp)
for _ in range(20): two()
pr = cProfile.Profile(); pr.enable()
for _ in range(300): two()
torch.cuda.synchronize(); pr.disable()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(30); print(s.getvalue()[:6000])
