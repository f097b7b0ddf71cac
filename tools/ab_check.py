"""Parity spot-check of the library currently installed as libnlbac_b200.so (used by tools/ab_bench.sh per variant):
TF32 tier vs the exact FP32 tier at a multi-chunk batch (the pipelined sweep + dW path) and at a small batch
(single chunk), flat max-norm of d-theta and L2 of dL/dy0.  Prints one line."""
import os
import sys

R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
import torch  # noqa: E402
import nlbac_b200  # noqa: E402

torch.manual_seed(0)
m = nlbac_b200.NeuralODEModel(3, 3, 6).cuda()
ps = list(m.parameters())
t20 = 0.02 * torch.arange(21, dtype=torch.float32)
out = []
for B in (int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 17, 1000):
    g = torch.Generator(device="cuda").manual_seed(7)
    y0 = torch.rand(B, 5, generator=g, device="cuda") * 2 - 1

    def grads(prec):
        yy = y0.clone().requires_grad_(True)
        sol = nlbac_b200.odeint(m, yy, t20, method="rk4", precision=prec)
        o = torch.autograd.grad(sol[-1][:, :3].sum(), [yy] + ps)
        return sol[-1].detach(), o[0], torch.cat([x.reshape(-1) for x in o[1:]])

    a, b = grads("tf32"), grads("fp32")
    rel = lambda x, y: ((x - y).abs().max() / y.abs().max()).item()  # noqa: E731
    l2 = lambda x, y: ((x - y).double().norm() / y.double().norm()).item()  # noqa: E731
    out.append("B=%d sol %.1e gy0(l2) %.1e dtheta(flat max) %.1e" % (B, rel(a[0], b[0]), l2(a[1], b[1]), rel(a[2], b[2])))
print("ab_check:", " | ".join(out))
