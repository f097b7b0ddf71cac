"""Oracle-side emulation of reduced-precision hidden-layer contractions (TEST INFRASTRUCTURE, see oracle/__init__.py).

Wraps the Linear/ReLU stacks of the reference NODE restatements so that the HIDDEN H x H contractions round their
operands like the tensor-core tiers do (TF32 round-to-nearest for the forward and dX GEMMs, bf16 for the dW GEMM,
products accumulated in higher precision), while the thin first / last layers stay fp32 — the arithmetic of
csrc/node_tc*.cuh restated in PyTorch.  tests/test_node_gpu.py holds the CUDA tier to THIS oracle at 2e-3 and to
the fp32 reference at the tier's declared bar; tools/prec_emul.py uses it to measure what each tier costs.
Mode strings: "fp32", "bf16", "tf32", "bf16x2", "tf32x3", "bf16x3", or "<fwd/dX mode>+<dW mode>".
"""
import copy

import torch
import torch.nn as nn


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
        return mm(x, W.t(), mode.split("+")[0]) + b
    @staticmethod
    def backward(ctx, g):
        x, W = ctx.saved_tensors
        m = ctx.mode.split("+")
        return mm(g, W, m[0]), mm(g.t(), x, m[-1]), g.sum(0), None

class EmuSeq(nn.Module):
    def __init__(self, seq, mode):
        super().__init__(); self.seq, self.mode = seq, mode
    def forward(self, x):
        lin = [m for m in self.seq if isinstance(m, nn.Linear)]
        for i, l in enumerate(lin):
            hidden = 0 < i < len(lin) - 1
            if hidden:
                # csrc/node_tc.cuh TcCfg::FOLD: where the K extent has a spare column (H % 8 != 0) the hidden bias is a
                # column of the weight image and is rounded with it (its gradient is unaffected)
                b = l.bias
                if self.mode.split("+")[0] == "tf32" and l.in_features % 8 != 0:
                    b = b + (rtf32(b.detach()) - b.detach())
                x = EmuLinear.apply(x, l.weight, b, self.mode)
            else:
                x = nn.functional.linear(x, l.weight, l.bias)
            if i < len(lin) - 1: x = torch.relu(x)
        return x



def emulate(model, mode="tf32+bf16"):
    """Deep copy of an oracle NODE (AffineNODE / BlackBoxNODE) whose hidden layers use `mode` arithmetic."""
    m = copy.deepcopy(model)
    m.zero_grad()
    if hasattr(m, "f_net"):
        m.f_net, m.g_net = EmuSeq(m.f_net, mode), EmuSeq(m.g_net, mode)
    else:
        m.net = EmuSeq(m.net, mode)
    return m
