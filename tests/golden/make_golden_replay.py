"""Golden vectors for the replay memory, produced by the REFERENCE's own ``ReplayMemory``
(NLBAC_Unicycle_RL_training/Unicycle_RL_training/sac_cbf_clf/replay_memory.py, imported by path in the build
container).  Run here:  ``python tests/golden/make_golden_replay.py``."""
import importlib.util
import os

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/NLBAC_Unicycle_RL_training/Unicycle_RL_training/sac_cbf_clf/replay_memory.py"


def transitions(n, with_t, seed):
    """Seeded float64 transitions shaped like the Unicycle env's (obs 7, action 2, scalars, centre positions 2)."""
    rng = np.random.default_rng(seed)
    out = []
    for k in range(n):
        tr = [rng.normal(size=7), rng.normal(size=2), float(rng.normal()), float(rng.random() < 0.2), rng.normal(size=2),
              rng.normal(size=2), rng.normal(size=7), float(rng.random() < 0.9)]
        tr += [np.array([0.02 * k]), np.array([0.02 * (k + 1)])] if with_t else [None, None]
        out.append(tuple(tr))
    return out


def case(with_t, seed, capacity=50, pushes=80, sizes=(16, 16, 50)):
    spec = importlib.util.spec_from_file_location("ref_replay", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    mem = mod.ReplayMemory(capacity, seed)
    samples = []
    trs = transitions(pushes, with_t, seed)
    for k, tr in enumerate(trs):
        mem.push(*tr)
        if k == 29:                                  # a sample before the ring is full
            samples.append(mem.sample(8))
    for b in sizes:
        samples.append(mem.sample(b))
    conv = lambda s: [None if a.dtype == object else torch.FloatTensor(a) for a in s]  # noqa: E731
    return {"with_t": with_t, "seed": seed, "capacity": capacity, "pushes": pushes, "sizes": [8] + list(sizes),
            "samples": [conv(s) for s in samples], "position": mem.position, "len": len(mem)}


if __name__ == "__main__":
    torch.save({"U": case(False, 11), "C": case(True, 12)}, os.path.join(HERE, "replay.pt"))
    print("replay ok")
