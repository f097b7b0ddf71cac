"""Forward / backward time of one precision tier at a given batch (CUDA events, 3 reps after warm-up).
   python tools/time_bwd.py <log2B> <precision> <params: 0|1>"""
import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
import torch, nlbac_b200
torch.manual_seed(0)
lb, prec, wantp = int(sys.argv[1]), sys.argv[2], int(sys.argv[3])
m = nlbac_b200.NeuralODEModel(3, 3, 6).cuda()
if not wantp:
    for p in m.parameters(): p.requires_grad_(False)
B = 1 << lb
y0 = (torch.rand(B, 5, device="cuda") * 2 - 1).requires_grad_(True)
t = 0.02 * torch.arange(21, dtype=torch.float32)
ev = lambda: torch.cuda.Event(enable_timing=True)
fw, bw = [], []
for i in range(4):
    e0, e1, e2 = ev(), ev(), ev()
    e0.record()
    loss = nlbac_b200.odeint(m, y0, t, method="rk4", precision=prec)[-1][:, :3].sum()
    e1.record()
    torch.autograd.grad(loss, [y0] + ([p for p in m.parameters()] if wantp else []))
    e2.record(); torch.cuda.synchronize()
    if i: fw.append(e0.elapsed_time(e1)); bw.append(e1.elapsed_time(e2))
print("B=2^%d %s params=%d  fwd %.1f ms  bwd %.1f ms" % (lb, prec, wantp, sum(fw)/3, sum(bw)/3))
