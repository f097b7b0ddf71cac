import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import torch, nlbac_b200
from test_node_gpu import _models, _y0, _rel
from oracle.torchdiffeq_restated import odeint as oracle_odeint
tag, B, method = "U", 40000, "euler"
ref, ours = _models(tag); y0 = _y0(tag, B); t = 0.02 * torch.arange(2, dtype=torch.float32)
ya = y0.clone().double().requires_grad_(True)
oracle_odeint(ref.double(), ya, t.double(), method=method).pow(2).sum().backward()
want = [p.grad.clone() for p in ref.parameters()]; ref.float()
res = {}
for prec in ("3xtf32", "fp32"):
    yb = y0.cuda().requires_grad_(True); ours.zero_grad()
    nlbac_b200.odeint(ours, yb, t, method=method, precision=prec).pow(2).sum().backward()
    res[prec] = [p.grad.clone() for p in ours.parameters()]
names = [n for n, _ in ours.named_parameters()]
for i, n in enumerate(names):
    print("%-16s 3x-vs-f64 %.1e  fp32-vs-f64 %.1e  3x-vs-fp32 %.1e  |g|max %.2e" % (n, _rel(res["3xtf32"][i], want[i]), _rel(res["fp32"][i], want[i]), _rel(res["3xtf32"][i], res["fp32"][i]), want[i].abs().max()))
