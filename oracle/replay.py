"""CPU restatement of the reference's replay memory (U sac_cbf_clf/replay_memory.py:4-28).

TEST INFRASTRUCTURE (see oracle/__init__.py).  Pinned against the reference's own class executed in the build
container (tests/golden/make_golden_replay.py -> tests/golden/replay.pt -> tests/test_oracle_replay.py)."""
import random

import numpy as np


class ReplayMemory:
    def __init__(self, capacity, seed):
        random.seed(seed)                                  # :7  (Python's GLOBAL generator)
        self.capacity, self.buffer, self.position = capacity, [], 0

    def push(self, *transition):                           # :12-18: ring of tuples, oldest overwritten
        if len(transition) in (8, 9):                      # 9: BU / BP with barrier_signal (BP replay_memory.py:12)
            transition = transition + (None, None)
        if len(self.buffer) < self.capacity:
            self.buffer.append(None)
        self.buffer[self.position] = tuple(transition)
        self.position = (self.position + 1) % self.capacity

    def batch_push(self, *batches):                        # P replay_memory.py:20-26: row-by-row push
        n = len(batches[0])
        for i in range(n):
            self.push(*[None if b is None else b[i] for b in batches])

    def sample(self, batch_size):                          # :21-25: random.sample of the tuples, np.stack per field
        batch = random.sample(self.buffer, batch_size)
        return tuple(map(np.stack, zip(*batch)))

    def __len__(self):
        return len(self.buffer)
