"""Cycle breakdown of the MMA round trip in pass A (debug build: NLBAC_NVCC_EXTRA=-DNLBAC_TC_PROF NLBAC_FORCE_BUILD=1).
Prints, for CTA 0 and each group's issuing thread: epilogue (previous done -> next layer entry, i.e. the owner
thread's own work + operand publication), wait for the weight packet, MMA issue + commits, wait for completion."""
import sys, os
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import torch, nlbac_b200
from test_node_gpu import _models, _y0
ref, ours = _models("U")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 148 * 3 * 128 * 4
y0 = _y0("U", B).cuda().requires_grad_(True)
t = 0.02 * torch.arange(21, dtype=torch.float32)
for _ in range(2):
    ours.zero_grad()
    nlbac_b200.odeint(ours, y0, t, method="rk4", precision="tf32")[-1].sum().backward()
    torch.cuda.synchronize()
