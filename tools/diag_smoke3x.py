"""Per-tensor error of the 3xTF32 tier on smoke()'s 512-row case, vs the fp64 oracle, the fp32 oracle and the FP32 tier."""
import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
import torch, nlbac_b200
from oracle import nlbac_ref
from oracle.torchdiffeq_restated import odeint as oracle_odeint
torch.manual_seed(0)
ref = nlbac_ref.AffineNODE(3, 3, 6)
ours = nlbac_b200.NeuralODEModel(3, 3, 6); ours.load_state_dict(ref.state_dict()); ours = ours.cuda()
_ = torch.randn(64, 5)
t = 0.02 * torch.arange(4, dtype=torch.float32)
y1 = torch.randn(512, 5)
def run_ref(dtype):
    m = nlbac_ref.AffineNODE(3, 3, 6); m.load_state_dict(ref.state_dict()); m = m.to(dtype)
    y = y1.to(dtype).detach().clone().requires_grad_(True)
    oracle_odeint(m, y, t.to(dtype), method="rk4")[-1].pow(2).sum().backward()
    return y.grad, [p.grad for p in m.parameters()]
def run_ours(prec):
    for p in ours.parameters(): p.grad = None
    y = y1.cuda().detach().requires_grad_(True)
    nlbac_b200.odeint(ours, y, t, method="rk4", precision=prec)[-1].pow(2).sum().backward()
    return y.grad.cpu(), [p.grad.cpu().clone() for p in ours.parameters()]
rel = lambda a, b: ((a.double() - b.double()).abs().max() / b.double().abs().max()).item()
g64, g32 = run_ref(torch.float64), run_ref(torch.float32)
names = [n for n, _ in ours.named_parameters()]
for prec in ("fp32", "3xtf32", "tf32"):
    gy, gp = run_ours(prec)
    print(prec, "dy0 vs f64 %.2e" % rel(gy, g64[0]), " row-wise worst rows:", ((gy.double() - g64[0]).abs().max(1).values / g64[0].abs().max()).topk(3).values.tolist())
    for n, a, b64, b32 in zip(names, gp, g64[1], g32[1]):
        print("   %-22s vs f64 %.2e  vs f32-oracle %.2e" % (n, rel(a, b64), rel(a, b32)))
res = {p: run_ours(p) for p in ("fp32", "3xtf32")}
i6 = names.index("f_net.6.bias")
o, a, b = g64[1][i6].double(), res["fp32"][1][i6].double(), res["3xtf32"][1][i6].double()
print("f_net.6.bias scale", o.abs().max().item())
for idx in ((a - o).abs().topk(3).indices.tolist() + (b - o).abs().topk(3).indices.tolist()):
    print(idx, "oracle %.9e fp32 %.9e 3x %.9e" % (o[idx].item(), a[idx].item(), b[idx].item()))
# run each tier twice: is it reproducible?
r2 = run_ours("fp32"); print("fp32 rerun diff", max(rel(x, y) for x, y in zip(r2[1], res["fp32"][1])))
r3 = run_ours("3xtf32"); print("3x rerun diff", max(rel(x, y) for x, y in zip(r3[1], res["3xtf32"][1])))
