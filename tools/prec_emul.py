"""Emulate reduced-precision hidden-layer contractions (bf16 / bf16x2 / bf16x3 / tf32 / 3xtf32, fp32 accumulate)
inside the oracle's NODE and report max-norm relative error of solution, dL/dy0, dL/dtheta vs fp64.
Test/analysis tooling only (imports oracle/)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn as nn
from oracle import nlbac_ref
from oracle.torchdiffeq_restated import odeint

def rbf16(x): return x.to(torch.bfloat16).to(torch.float32)
def rtf32(x):
    i = x.contiguous().view(torch.int32)
    i = (i + 0x1000) & ~0x1FFF      # round-to-nearest (ties away) to 10 mantissa bits
    return i.view(torch.float32)
def split(x, r, n):
    parts, rem = [], x
    for _ in range(n):
        p = r(rem); parts.append(p); rem = rem - p
    return parts
def mm(a, b, mode):
    """a @ b with emulated operand precision; products accumulate in fp32 (emulated via fp64 then cast)."""
    if mode == "fp32": return a @ b
    r, n = {"bf16": (rbf16, 1), "bf16x2": (rbf16, 2), "bf16x3": (rbf16, 3), "tf32": (rtf32, 1), "tf32x3": (rtf32, 2)}[mode]
    A, B = split(a, r, n), split(b, r, n)
    out = 0
    if mode in ("bf16", "tf32"): return (A[0].double() @ B[0].double()).float()
    if mode == "bf16x2" or mode == "tf32x3":   # hi*hi + hi*lo + lo*hi
        return (A[0].double() @ B[0].double() + A[0].double() @ B[1].double() + A[1].double() @ B[0].double()).float()
    # bf16x3: 6 products up to order 2
    acc = 0
    for i in range(3):
        for j in range(3):
            if i + j <= 2: acc = acc + A[i].double() @ B[j].double()
    return acc.float()

class EmuLinear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, W, b, mode):
        ctx.save_for_backward(x, W); ctx.mode = mode
        return mm(x, W.t(), mode) + b
    @staticmethod
    def backward(ctx, g):
        x, W = ctx.saved_tensors
        return mm(g, W, ctx.mode), mm(g.t(), x, ctx.mode), g.sum(0), None

class EmuSeq(nn.Module):
    def __init__(self, seq, mode):
        super().__init__(); self.seq, self.mode = seq, mode
    def forward(self, x):
        lin = [m for m in self.seq if isinstance(m, nn.Linear)]
        for i, l in enumerate(lin):
            hidden = 0 < i < len(lin) - 1
            x = EmuLinear.apply(x, l.weight, l.bias, self.mode) if hidden else nn.functional.linear(x, l.weight, l.bias)
            if i < len(lin) - 1: x = torch.relu(x)
        return x

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
    for mode in ["fp32", "bf16", "tf32", "bf16x2", "tf32x3", "bf16x3"]:
        r = run(tag, mode)
        print("%-7s sol %.2e  gy0 %.2e  gtheta(max over tensors) %.2e" % (
            mode, rel(r[0], ref[0]), rel(r[1], ref[1]), max(rel(a, b) for a, b in zip(r[2], ref[2]))))
