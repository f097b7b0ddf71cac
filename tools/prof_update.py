"""Kernel list of ONE GraphedUpdate replay (Unicycle, B = 128) via torch.profiler: count, total and mean time per kernel name."""
import os, sys, collections
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import numpy as np, torch, nlbac_b200 as nb
from test_update_gpu import _build, _Env, SHAPES
C = nb.constraints
g = torch.load(os.path.join(R, "tests", "golden", "update_U.pt"), weights_only=False)
k, B = g["consts"], g["B"]
nets = _build(g, nb, "U"); node, lya = nets["neural_ode_model"], nets["lyapunovNet"]
kw = dict(dt=k["dt"], gamma_b=k["gamma_b"], hazards_locations=k["hazards"], hazards_radius=k["hazards_radius"], l_p=k["l_p"], batch_size=B)
upd = nb.GraphedUpdate(nets["critic"], nets["critic_target"], lya, nets["lyapunovNet_target"], nets["policy"], nets["backup_policy"],
                       node, nb.DynamicsModel(_Env("Unicycle")), B,
                       primary=lambda s, pi, V: C.unicycle_required_matrix(node, lya, s, pi, V, **kw),
                       backup=lambda s, pi: C.unicycle_required_matrix(node, None, s, pi, None, **kw),
                       num_constraints=8, obs_dim=7, act_dim=2, center_pos_num=2)
batch = tuple(None if x is None else x.cuda() for x in g["batches"][0])
for _ in range(3): upd.step(batch)
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    upd.graph.replay(); torch.cuda.synchronize()
agg = collections.OrderedDict()
for e in prof.events():
    if e.device_type == torch.autograd.DeviceType.CUDA:
        a = agg.setdefault(e.name[:70], [0, 0.0]); a[0] += 1; a[1] += e.device_time
tot = sum(v[1] for v in agg.values()); n = sum(v[0] for v in agg.values())
print("kernels %d  sum of kernel time %.0f us" % (n, tot))
for name, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1])[:28]:
    print("%4d x %6.1f us = %7.1f us  %s" % (c, t / c, t, name))
