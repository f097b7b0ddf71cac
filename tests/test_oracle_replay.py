"""Pin oracle/replay.py to the golden vectors produced by the REFERENCE's own ReplayMemory
(tests/golden/make_golden_replay.py, run in the build container)."""
import os
import sys

import pytest
import torch

from oracle.replay import ReplayMemory

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
from make_golden_replay import transitions  # noqa: E402


def replay_sequence(mem, g):
    """The push / sample sequence the golden run recorded; returns the sampled batches."""
    import numpy as np
    out = []
    trs = transitions(g["pushes"], g["with_t"], g["seed"], g.get("barrier", False))
    if g.get("batch"):
        cols = lambda rows: [np.stack([np.asarray(r[j]) for r in rows]) for j in range(8)]  # noqa: E731
        mem.batch_push(*cols(trs[:30]))
        out.append(mem.sample(8))
        mem.batch_push(*cols(trs[30:]))
    else:
        for k, tr in enumerate(trs):
            mem.push(*tr)
            if k == 29:
                out.append(mem.sample(8))
    for b in g["sizes"][1:]:
        out.append(mem.sample(b))
    return out


@pytest.mark.parametrize("tag", ["U", "C", "BP", "BPt", "Pbatch"])
def test_oracle_replay_matches_reference(tag, golden_dir):
    g = torch.load(os.path.join(golden_dir, "replay.pt"), weights_only=False)[tag]
    mem = ReplayMemory(g["capacity"], g["seed"])
    got = replay_sequence(mem, g)
    assert mem.position == g["position"] and len(mem) == g["len"]
    for s, w in zip(got, g["samples"]):
        for a, b in zip(s, w):
            if b is None:
                assert a.dtype == object
            else:
                assert torch.equal(torch.FloatTensor(a), b)
