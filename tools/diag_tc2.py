"""Per-parameter errors of the TF32 tier on the sub-stepping / multi-output parity case (tests/test_node_gpu.py)."""
import sys, os
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import torch, nlbac_b200
from test_node_gpu import _models, _y0, _rel
from oracle.torchdiffeq_restated import odeint as oracle_odeint
from oracle.tf32_emul import emulate
for method, step_size, B in [("euler", None, 4500), ("euler", None, 4501), ("midpoint", None, 4500), ("rk4", 0.015, 4500), ("euler", 0.007, 4500)]:
    ref, ours = _models("U")
    emu = emulate(ref, "tf32+bf16")
    torch.manual_seed(B)
    y0 = _y0("U", B)
    t = torch.tensor([0.0, 0.02, 0.05, 0.06])
    opts = {} if step_size is None else {"options": {"step_size": step_size}}
    w = torch.linspace(0.5, 1.5, 4).reshape(4, 1, 1)
    res = {}
    for name, m in (("fp32", ref), ("emu", emu)):
        m.zero_grad()
        ya = y0.clone().requires_grad_(True)
        (oracle_odeint(m, ya, t, method=method, **opts) * w).pow(2).sum().backward()
        res[name] = [ya.grad] + [p.grad.clone() for p in m.parameters()]
    ours.zero_grad()
    yb = y0.cuda().requires_grad_(True)
    (nlbac_b200.odeint(ours, yb, t, method=method, precision="tf32", **opts) * w.cuda()).pow(2).sum().backward()
    got = [yb.grad] + [p.grad for p in ours.parameters()]
    names = ["y0"] + [k for k, _ in ours.named_parameters()]
    print("==", method, step_size, B)
    for n, g, a, e in zip(names, got, res["fp32"], res["emu"]):
        print("  %-16s ours-vs-fp32 %.2e   emu-vs-fp32 %.2e   ours-vs-emu %.2e" % (n, _rel(g, a), _rel(e, a), _rel(g, e)))
