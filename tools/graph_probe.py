"""Does a launch-bound NODE forward + backward capture into a CUDA graph, and what does a replay cost?"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, nlbac_b200
m = nlbac_b200.NeuralODEModel(3, 3, 6).cuda()
ps = list(m.parameters())
y = torch.rand(128, 5, device="cuda")
t = [0.0, 0.02]
def step():
    sol = nlbac_b200.odeint(m, y, t, method="euler")
    return torch.autograd.grad(sol[-1][:, :3].sum(), ps)
s = torch.cuda.Stream()
s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    for _ in range(3): ref = step()
torch.cuda.current_stream().wait_stream(s)
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
try:
    with torch.cuda.graph(g):
        out = step()
    g.replay(); torch.cuda.synchronize()
    err = max(((a - b).abs().max() / b.abs().max()).item() for a, b in zip(out, ref))
    print("captured; replay vs eager rel err %.2e" % err)
    y.copy_(torch.rand(128, 5, device="cuda")); g.replay(); torch.cuda.synchronize()
    ref2 = step(); torch.cuda.synchronize()
    print("new input rel err %.2e" % max(((a - b).abs().max() / b.abs().max()).item() for a, b in zip(out, ref2)))
    for _ in range(20): g.replay()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(300): g.replay()
    torch.cuda.synchronize(); print("graph replay %.1f us per fwd+bwd" % ((time.perf_counter() - t0) / 300 * 1e6))
except Exception as e:
    print("capture failed:", repr(e)[:400])
