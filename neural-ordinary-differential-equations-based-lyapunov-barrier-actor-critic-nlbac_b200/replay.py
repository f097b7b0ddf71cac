"""Device-resident replay memory with the reference's interface (U sac_cbf_clf/replay_memory.py:4-28).

``ReplayMemory(capacity, seed)``, ``push(state, action, reward, constraint, center_pos, next_center_pos, next_state,
mask, t=None, next_t=None)``, ``sample(batch_size)``, ``len()``, ``.position`` behave like the reference's class: the
ring overwrites the oldest transition, ``random.seed(seed)`` seeds Python's global generator and ``sample`` draws the
SAME indices ``random.sample(self.buffer, batch_size)`` would (``random.sample`` depends on the population's length
only).  What changes is where the data lives: each transition is one float32 row of a ``[capacity, width]`` device
tensor (rows staged in pinned host memory and flushed with one copy per contiguous range), and a minibatch is ONE
gather launch (``nlbac_replay_gather``) instead of ten ``np.stack`` + ten H2D copies per update
(U sac_cbf_clf.py:201-229).  ``sample`` returns float32 CUDA tensors shaped like the reference's stacked arrays
(what ``torch.FloatTensor(x).to(device)`` would have produced); ``sample(batch_size, as_numpy=True)`` returns host
arrays for unmodified callers.
"""
import random

import numpy as np
import torch

from . import _lib

FIELDS = ("state", "action", "reward", "constraint", "center_pos", "next_center_pos", "next_state", "mask", "t",
          "next_t")


class ReplayMemory:
    def __init__(self, capacity, seed, device="cuda"):
        random.seed(seed)
        self.capacity = int(capacity)
        self.position = 0
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("nlbac_b200.ReplayMemory lives on a CUDA device (no CPU fallback)")
        self._n = 0
        self._layout = None      # [(name, offset, width, shape or None)]
        self._width = 0
        self._ring = None        # device [capacity, width]
        self._stage = None       # pinned host [capacity, width]
        self._dirty = []         # row indices written since the last flush, in push order

    def __len__(self):
        return self._n

    def _init_layout(self, fields):
        off, lay = 0, []
        for name, v in zip(FIELDS, fields):
            if v is None:
                lay.append((name, off, 0, None))
                continue
            a = np.asarray(v, dtype=np.float32)
            lay.append((name, off, a.size, a.shape))
            off += a.size
        self._layout, self._width = lay, off
        self._ring = torch.empty((self.capacity, off), dtype=torch.float32, device=self.device)
        self._stage = torch.empty((self.capacity, off), dtype=torch.float32).pin_memory()

    def push(self, state, action, reward, constraint, center_pos, next_center_pos, next_state, mask, t=None,
             next_t=None):
        fields = (state, action, reward, constraint, center_pos, next_center_pos, next_state, mask, t, next_t)
        if self._layout is None:
            self._init_layout(fields)
        row = self._stage[self.position].numpy()
        for (name, off, w, shape), v in zip(self._layout, fields):
            if (v is None) != (w == 0):
                raise ValueError("ReplayMemory.push: field %r changed between None and a value" % name)
            if w:
                a = np.asarray(v, dtype=np.float32).reshape(-1)
                if a.size != w:
                    raise ValueError("ReplayMemory.push: field %r has %d values, expected %d" % (name, a.size, w))
                row[off:off + w] = a
        self._dirty.append(self.position)
        if self._n < self.capacity:
            self._n += 1
        self.position = (self.position + 1) % self.capacity

    def _flush(self):
        if not self._dirty:
            return
        rows = sorted(set(self._dirty))
        self._dirty = []
        start = prev = rows[0]
        for r in rows[1:] + [None]:
            if r is not None and r == prev + 1:
                prev = r
                continue
            self._ring[start:prev + 1].copy_(self._stage[start:prev + 1], non_blocking=True)
            if r is not None:
                start = prev = r

    def sample(self, batch_size, as_numpy=False):
        if batch_size > self._n:
            raise ValueError("Sample larger than population or is negative")     # random.sample's own error
        idx = random.sample(range(self._n), batch_size)
        self._flush()
        idx_dev = torch.tensor(idx, dtype=torch.int64).pin_memory().to(self.device, non_blocking=True)
        out = torch.empty((batch_size, self._width), dtype=torch.float32, device=self.device)
        _lib.check(_lib.lib().nlbac_replay_gather(self._ring.data_ptr(), self._width, idx_dev.data_ptr(), batch_size,
                                                  out.data_ptr(), _lib.raw_stream()), "replay_gather")
        res = []
        for name, off, w, shape in self._layout:
            if w == 0:
                res.append(None)
                continue
            col = out[:, off:off + w]
            res.append(col.reshape((batch_size,) + tuple(shape)))
        if as_numpy:
            res = [None if r is None else r.cpu().numpy() for r in res]
        return tuple(res)
