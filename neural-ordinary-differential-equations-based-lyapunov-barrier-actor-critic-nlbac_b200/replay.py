"""Device-resident replay memory with the reference's interface (U sac_cbf_clf/replay_memory.py:4-28).

``ReplayMemory(capacity, seed)``, ``push(state, action, reward, constraint, center_pos, next_center_pos, next_state,
mask, t=None, next_t=None)`` -- or the 11-field form of the learned-barrier variants with ``barrier_signal`` after
``constraint`` (BU / BP replay_memory.py:12); the field count of the first push fixes the layout --, ``batch_push``
(P / BP :20-26), ``sample(batch_size)``, ``len()``, ``.position`` behave like the reference's class: the
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
# learned-barrier variants (BU / BP sac_cbf_clf/replay_memory.py:12): an extra ``barrier_signal`` after ``constraint``
FIELDS_BARRIER = ("state", "action", "reward", "constraint", "barrier_signal", "lya_pre_term", "lya_next_term",
                  "next_state", "mask", "t", "next_t")


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
        names = FIELDS_BARRIER if len(fields) == len(FIELDS_BARRIER) else FIELDS
        off, lay = 0, []
        for name, v in zip(names, fields):
            if v is None:
                lay.append((name, off, 0, None))
                continue
            a = np.asarray(v, dtype=np.float32)
            lay.append((name, off, a.size, a.shape))
            off += a.size
        self._layout, self._width = lay, off
        self._ring = torch.empty((self.capacity, off), dtype=torch.float32, device=self.device)
        self._stage = torch.empty((self.capacity, off), dtype=torch.float32).pin_memory()

    def _fields(self, args, t, next_t):
        """The reference's positional signatures: 8 (+ t, next_t) fields for U / C / P, 9 (+ t, next_t) with
        ``barrier_signal`` for BU / BP; ``t`` / ``next_t`` may also come as keywords."""
        n = len(args)
        if n in (10, 11):
            if t is not None or next_t is not None:
                raise TypeError("ReplayMemory.push: t / next_t given twice")
            return tuple(args)
        if n in (8, 9):
            return tuple(args) + (t, next_t)
        raise TypeError("ReplayMemory.push takes 8 or 9 transition fields (+ t, next_t), got %d" % n)

    def push(self, *args, t=None, next_t=None):
        fields = self._fields(args, t, next_t)
        if self._layout is None:
            self._init_layout(fields)
        if len(fields) != len(self._layout):
            raise TypeError("ReplayMemory.push: %d fields, the memory was started with %d" % (len(fields),
                                                                                               len(self._layout)))
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

    def batch_push(self, *batches, t_batch=None, next_t_batch=None):
        """P / BP ``batch_push`` (P sac_cbf_clf/replay_memory.py:20-26): the rows of the batches, pushed in order --
        here as whole-column copies into the staging ring instead of a Python loop over ``push``."""
        fields = self._fields(batches, t_batch, next_t_batch)
        n = len(np.asarray(fields[0]))
        if n == 0:
            return
        if self._layout is None:
            self._init_layout(tuple(None if f is None else np.asarray(f)[0] for f in fields))
        if len(fields) != len(self._layout):
            raise TypeError("ReplayMemory.batch_push: %d fields, the memory was started with %d" % (
                len(fields), len(self._layout)))
        idx = (self.position + np.arange(n)) % self.capacity
        keep = slice(max(0, n - self.capacity), n)          # more rows than capacity: only the last `capacity` survive
        stage = self._stage.numpy()
        for (name, off, w, shape), v in zip(self._layout, fields):
            if (v is None) != (w == 0):
                raise ValueError("ReplayMemory.batch_push: field %r changed between None and a value" % name)
            if w:
                a = np.asarray(v, dtype=np.float32).reshape(n, -1)
                if a.shape[1] != w:
                    raise ValueError("ReplayMemory.batch_push: field %r has %d values per row, expected %d" % (
                        name, a.shape[1], w))
                stage[idx[keep], off:off + w] = a[keep]
        self._dirty.extend(int(i) for i in idx[keep])
        self._n = min(self.capacity, self._n + n)
        self.position = int((self.position + n) % self.capacity)

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

    def sample_into(self, batch_size, outs):
        """``sample`` straight into caller-owned contiguous float32 CUDA buffers (``outs``: one per field in layout
        order, ``None`` or a missing tail = skip): the same indices ``sample`` would draw, ONE launch, no per-field
        copies.  ``GraphedUpdate.update_parameters`` uses it to fill the static inputs of the captured update."""
        import ctypes
        if batch_size > self._n:
            raise ValueError("Sample larger than population or is negative")
        idx = random.sample(range(self._n), batch_size)
        self._flush()
        idx_dev = torch.tensor(idx, dtype=torch.int64).pin_memory().to(self.device, non_blocking=True)
        lay = [(off, w) for _, off, w, _ in self._layout if w > 0]
        fields = [i for i, (_, _, w, _) in enumerate(self._layout) if w > 0]
        n = len(lay)
        ptrs = []
        for k, fi in enumerate(fields):
            o = outs[fi] if fi < len(outs) else None
            if o is not None:
                if not (o.is_cuda and o.dtype == torch.float32 and o.is_contiguous() and o.numel() == batch_size * lay[k][1]):
                    raise ValueError("sample_into: buffer %d must be contiguous float32 CUDA with %d x %d elements" % (
                        fi, batch_size, lay[k][1]))
            ptrs.append(o.data_ptr() if o is not None else None)
        _lib.check(_lib.lib().nlbac_replay_gather_split(
            self._ring.data_ptr(), self._width, idx_dev.data_ptr(), batch_size, n,
            (ctypes.c_int * n)(*[o for o, _ in lay]), (ctypes.c_int * n)(*[w for _, w in lay]),
            (ctypes.c_void_p * n)(*ptrs), _lib.raw_stream()), "replay_gather_split")

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
