"""Emulate reduced-precision hidden-layer contractions (bf16 / bf16x2 / bf16x3 / tf32 / 3xtf32, fp32 accumulate)
inside the oracle's NODE and report max-norm relative error of solution, dL/dy0, dL/dtheta vs fp64.
Test/analysis tooling only (imports oracle/)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn as nn
from oracle import nlbac_ref
from oracle.torchdiffeq_restated import odeint

from oracle.tf32_emul import EmuSeq, mm, rbf16, rtf32  # noqa: E402,F401

def run(tag, mode, B=2048, steps=20, dtype=torch.float32):
    torch.manual_seed(0)
    m = nlbac_ref.AffineNODE(3, 3, 6) if tag == "U" else nlbac_ref.AffineNODE(6, 6, 12)
    m = m.to(dtype)
    if mode != "ref":
        m2 = nlbac_ref.AffineNODE(m.input_dim, m.output_dim1, m.output_dim2); m2.load_state_dict(m.state_dict())
        f, g = m2.f_net, m2.g_net
        m2.f_net, m2.g_net = EmuSeq(f, mode), EmuSeq(g, mode)
        m = m2
    g = torch.Generator().manual_seed(1)
    y0 = (torch.rand(B, m.n_s + m.n_u, generator=g) * 2 - 1).to(dtype).requires_grad_(True)
    t = 0.02 * torch.arange(steps + 1, dtype=dtype)
    sol = odeint(m, y0, t, method="rk4")
    sol[-1][:, :m.n_s].sum().backward()
    return sol[-1].detach().double(), y0.grad.double(), [p.grad.double() for p in m.parameters()]

def rel(a, b): return ((a - b).abs().max() / b.abs().max()).item()
if __name__ == "__main__":
    tag = sys.argv[1] if len(sys.argv) > 1 else "U"
    ref = run(tag, "ref", dtype=torch.float64)
    for mode in (sys.argv[2:] or ["fp32", "bf16", "tf32", "bf16x2", "tf32x3", "bf16x3"]):
        r = run(tag, mode)
        print("%-7s sol %.2e  gy0 %.2e  gtheta(max over tensors) %.2e" % (
            mode, rel(r[0], ref[0]), rel(r[1], ref[1]), max(rel(a, b) for a, b in zip(r[2], ref[2]))))
