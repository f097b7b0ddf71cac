import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
import test_dopri5_gpu as T
from test_node_gpu import _rel
for tag, B, t in [("U", 8, [0.0, 0.02]), ("U", 8, [0.0, 0.1]), ("U", 8, [0.0, 0.3]), ("U", 8, [0.0, 0.3, 0.5]), ("U", 8, [0.0, 1.0])]:
    got, want, st, st_ref = T._pair(tag, B, torch.tensor(t))
    print(t, "acc/rej", st["n_accepted"], st["n_rejected"], "ref", st_ref["n_accepted"], st_ref["n_rejected"],
          "sol %.1e gy0 %.1e gp %.1e" % (_rel(got[0], want[0]), _rel(got[1], want[1]), max(_rel(a, b) for a, b in zip(got[2], want[2]))))
    d = (got[1].cpu() - want[1]).abs()
    print("   err by column", d.max(0).values.tolist())
