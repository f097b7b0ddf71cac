import sys, os
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import torch, nlbac_b200
from test_node_gpu import _models, _y0, _rel
from oracle.torchdiffeq_restated import odeint as oracle_odeint
from tools.prec_emul import EmuSeq
for tag, B, nst in [("U", 256, 3), ("U", 4096, 3), ("U", 4096, 20), ("C", 256, 3), ("P", 256, 3)]:
    ref, ours = _models(tag)
    y0 = _y0(tag, B)
    t = 0.02 * torch.arange(nst + 1, dtype=torch.float32)
    ya = y0.clone().requires_grad_(True)
    oracle_odeint(ref, ya, t, method="rk4")[-1].pow(2).sum().backward()
    want = [ya.grad] + [p.grad.clone() for p in ref.parameters()]
    # emulated tf32 oracle (all hidden contractions tf32)
    import copy
    emu = copy.deepcopy(ref); emu.zero_grad()
    if hasattr(emu, "f_net"):
        emu.f_net, emu.g_net = EmuSeq(emu.f_net, "tf32+bf16"), EmuSeq(emu.g_net, "tf32+bf16")
    else:
        emu.net = EmuSeq(emu.net, "tf32+bf16")
    ye = y0.clone().requires_grad_(True)
    oracle_odeint(emu, ye, t, method="rk4")[-1].pow(2).sum().backward()
    em = [ye.grad] + [p.grad.clone() for p in emu.parameters()]
    ours.zero_grad()
    yb = y0.cuda().requires_grad_(True)
    nlbac_b200.odeint(ours, yb, t, method="rk4", precision="tf32")[-1].pow(2).sum().backward()
    got = [yb.grad] + [p.grad for p in ours.parameters()]
    names = ["y0"] + [k for k, _ in ours.named_parameters()]
    print("==", tag, B, nst)
    for n, g, w, e in zip(names, got, want, em):
        print("  %-16s ours-vs-fp32 %.2e   emu(tf32+bf16)-vs-fp32 %.2e   ours-vs-emu %.2e" % (n, _rel(g, w), _rel(e, w), _rel(g, e)))
