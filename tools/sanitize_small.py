"""Small-batch pass over the FP32 NODE kernels, the MLPs and the tail kernels for compute-sanitizer (racecheck / memcheck):
   compute-sanitizer --tool racecheck python tools/sanitize_small.py [tc]      (tc: also the tcgen05 TF32 tier)"""
import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
import torch, nlbac_b200 as nb
torch.manual_seed(0)
dev = "cuda"
tiers = ["fp32"] + (["tf32", "3xtf32"] if "tc" in sys.argv[1:] else [])
for mk, D in ((lambda: nb.NeuralODEModel(3, 3, 6), 5), (lambda: nb.NeuralODEModel(12, 10), 12)):
    m = mk().to(dev)
    for prec in tiers:
        for B, method, steps in ((33, "euler", 1), (200, "rk4", 2)):
            y0 = torch.randn(B, D, device=dev, requires_grad=True)
            t = 0.02 * torch.arange(steps + 1, dtype=torch.float32)
            nb.odeint(m, y0, t, method=method, precision=prec)[-1].sum().backward()
q, pol = nb.QNetwork(7, 2, 256).to(dev), nb.GaussianPolicy(7, 2, 256).to(dev)
for B in (5, 130, 5000):
    obs, act = torch.randn(B, 7, device=dev), torch.randn(B, 2, device=dev, requires_grad=True)
    q1, q2 = q(obs, act); (q1.sum() + q2.sum()).backward()
    a, lp, _ = pol.sample(obs, torch.randn(B, 2, device=dev)); (a.sum() + lp.sum()).backward()
    hz = [[0., 0.], [0., 1.5], [0., -1.5]]
    desc = nb.make_constraint_desc("unicycle", dt=0.02, gamma_b=50.0, gamma_l=1.0, batch_size=B, collision_radius=0.525, hazards=hz)
    psn, Vn = torch.randn(B, 2, device=dev, requires_grad=True), torch.rand(B, 1, device=dev, requires_grad=True)
    nb.constraint_sums(desc, torch.randn(B, 2, device=dev), psn, None, torch.rand(B, 1, device=dev), Vn).sum().backward()
    s3 = torch.randn(B, 3, device=dev, requires_grad=True); nb.lookahead(s3, 0.03).sum().backward()
opt = nb.FusedAdam(q.parameters(), lr=1e-3); opt.step()
nb.soft_update(nb.QNetwork(7, 2, 256).to(dev), q, 0.005)
torch.cuda.synchronize()
print("ok", tiers)
