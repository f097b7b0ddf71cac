"""Small TF32-tier forward + backward for ncu captures: one full pass per CTA (148 x 3 tiles of 128 rows) x 20 rk4 steps."""
import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
import torch, nlbac_b200
torch.manual_seed(0)
m = nlbac_b200.NeuralODEModel(3, 3, 6).cuda()
B = int(sys.argv[1]) if len(sys.argv) > 1 else 148 * 3 * 128
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
prec = sys.argv[3] if len(sys.argv) > 3 else "tf32"
y0 = (torch.rand(B, 5, device="cuda") * 2 - 1).requires_grad_(True)
t = 0.02 * torch.arange(steps + 1, dtype=torch.float32)
for _ in range(2):
    m.zero_grad()
    nlbac_b200.odeint(m, y0, t, method="rk4", precision=prec)[-1][:, :3].sum().backward()
    torch.cuda.synchronize()
print("ok", B, steps, prec)
