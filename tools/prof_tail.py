"""One launch of each streaming kernel (K3-K6, optimiser tail, env steps) at B = 2^20 rows for an ncu capture:
   ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,dram__throughput.avg.pct_of_peak_sustained_elapsed python tools/prof_tail.py"""
import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
import torch, nlbac_b200 as nb
torch.manual_seed(0)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
dev = "cuda"
q, lya = nb.QNetwork(7, 2, 256).to(dev), nb.LyaNetwork(2, 256).to(dev)
pol = nb.GaussianPolicy(7, 2, 256).to(dev)
obs, act = torch.randn(B, 7, device=dev), torch.randn(B, 2, device=dev, requires_grad=True)
for rep in range(2):
    q1, q2 = q(obs, act)
    (q1.sum() + q2.sum()).backward()
    a, lp, _ = pol.sample(obs, torch.randn(B, 2, device=dev))
    (a.sum() + lp.sum()).backward()
    env = type("E", (), {"dynamics_mode": "Pvtol"})
    dm = nb.DynamicsModel(env)
    st7 = torch.randn(B, 7, device=dev, requires_grad=True)
    o = dm.get_obs(st7); o.sum().backward()
    dm.get_state(o.detach())
    s3 = torch.randn(B, 3, device=dev, requires_grad=True)
    p = nb.lookahead(s3, 0.03); p.sum().backward()
    hz = [[0., 0.], [0., 1.5], [0., -1.5], [-1.5, 1.5], [-1.5, -1.5], [1.5, -1.5], [1.5, 1.5]]
    desc = nb.make_constraint_desc("unicycle", dt=0.02, gamma_b=50.0, gamma_l=1.0, batch_size=B, collision_radius=0.525, hazards=hz)
    ps, psn = torch.randn(B, 2, device=dev), torch.randn(B, 2, device=dev, requires_grad=True)
    V, Vn = torch.rand(B, 1, device=dev), torch.rand(B, 1, device=dev, requires_grad=True)
    nb.constraint_sums(desc, ps, psn, None, V, Vn).sum().backward()
    big = torch.nn.Linear(4096, 4096).to(dev); big2 = torch.nn.Linear(4096, 4096).to(dev)
    for p_ in big.parameters(): p_.grad = torch.randn_like(p_)
    nb.FusedAdam(big.parameters(), lr=1e-3).step()
    nb.soft_update(big2, big, 0.005)
    for E in (nb.VecUnicycleEnv(B), nb.VecSimulatedCarsEnv(B), nb.VecPvtolEnv(B)):
        n_u = 1 if isinstance(E, nb.VecSimulatedCarsEnv) else 2
        a_ = torch.rand(B, dtype=torch.float64, device=dev) * 2 - 1 if n_u == 1 else torch.rand(B, 2, dtype=torch.float64, device=dev) * 2 - 1
        E.step(a_)
    torch.cuda.synchronize()
print("ok", B)
