"""Golden vectors for the replay memory, produced by the REFERENCE's own ``ReplayMemory``
(NLBAC_Unicycle_RL_training/Unicycle_RL_training/sac_cbf_clf/replay_memory.py, imported by path in the build
container).  Run here:  ``python tests/golden/make_golden_replay.py``."""
import importlib.util
import os

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/NLBAC_Unicycle_RL_training/Unicycle_RL_training/sac_cbf_clf/replay_memory.py"
REF_P = "/root/reference/NLBAC_pvtol_RL_training/Pvtol_RL_training/sac_cbf_clf/replay_memory.py"
REF_BP = ("/root/reference/neural_barrier_certificate/neural_barrier_certificate_NLBAC_pvtol_RL_training/"
          "Pvtol_RL_training/sac_cbf_clf/replay_memory.py")


def transitions(n, with_t, seed, barrier=False):
    """Seeded float64 transitions shaped like the Unicycle env's (obs 7, action 2, scalars, centre positions 2);
    ``barrier``: the 9-field form of BU / BP with ``barrier_signal`` after ``constraint``."""
    rng = np.random.default_rng(seed)
    out = []
    for k in range(n):
        tr = [rng.normal(size=7), rng.normal(size=2), float(rng.normal()), float(rng.random() < 0.2), rng.normal(size=2),
              rng.normal(size=2), rng.normal(size=7), float(rng.random() < 0.9)]
        if barrier:
            tr.insert(4, -0.1 * float(rng.random() < 0.3))
        tr += [np.array([0.02 * k]), np.array([0.02 * (k + 1)])] if with_t else [None, None]
        out.append(tuple(tr))
    return out


def case(with_t, seed, capacity=50, pushes=80, sizes=(16, 16, 50), ref=REF, barrier=False, batch=False):
    spec = importlib.util.spec_from_file_location("ref_replay", ref)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    mem = mod.ReplayMemory(capacity, seed)
    samples = []
    trs = transitions(pushes, with_t, seed, barrier)
    if batch:                                        # P batch_push: 30 rows, a sample, then the rest in one call
        cols = lambda rows: [None if rows[0][j] is None else np.stack([np.asarray(r[j]) for r in rows])  # noqa: E731
                             for j in range(len(rows[0]))]
        mem.batch_push(*cols(trs[:30])[:8])
        samples.append(mem.sample(8))
        mem.batch_push(*cols(trs[30:])[:8])
    else:
        for k, tr in enumerate(trs):
            mem.push(*tr)
            if k == 29:                              # a sample before the ring is full
                samples.append(mem.sample(8))
    for b in sizes:
        samples.append(mem.sample(b))
    conv = lambda s: [None if a.dtype == object else torch.FloatTensor(a) for a in s]  # noqa: E731
    return {"barrier": barrier, "batch": batch, "with_t": with_t, "seed": seed, "capacity": capacity, "pushes": pushes, "sizes": [8] + list(sizes),
            "samples": [conv(s) for s in samples], "position": mem.position, "len": len(mem)}


if __name__ == "__main__":
    torch.save({"U": case(False, 11), "C": case(True, 12), "BP": case(False, 13, ref=REF_BP, barrier=True),
                "BPt": case(True, 14, ref=REF_BP, barrier=True), "Pbatch": case(False, 15, ref=REF_P, batch=True)},
               os.path.join(HERE, "replay.pt"))
    print("replay ok")
