"""A few fused-MLP forward + backward calls at B = 128 (for ncu)."""
import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
import torch, nlbac_b200 as nb
net = nb.LyaNetwork(11, 256).cuda()
x = torch.randn(128, 11, device="cuda", requires_grad=True)
for _ in range(4):
    y = net(x)
    torch.autograd.grad(y.sum(), [x] + list(net.parameters()))
torch.cuda.synchronize()
