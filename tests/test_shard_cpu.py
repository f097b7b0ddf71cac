"""World-size-2 gloo test of the trajectory sharding used by bench.py --gpus N: contiguous row
blocks per rank, replicated weights, ONE all-reduce (sum) of the flat d-theta buffer.  Runs on CPU with the oracle
standing in for the per-rank solve — what is under test is the host-side split / reduce logic, which must
reproduce the single-rank gradient."""
import os
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def shard_rows(n, world, rank):
    """Contiguous block of rows for `rank` (same rule as bench.py: equal blocks, remainder to the last ranks)."""
    base, rem = divmod(n, world)
    start = rank * base + max(0, rank - (world - rem))
    return start, start + base + (1 if rank >= world - rem else 0)


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    from oracle import nlbac_ref
    from oracle.torchdiffeq_restated import odeint
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(0)
    model = nlbac_ref.AffineNODE(3, 3, 6)
    y0 = torch.randn(37, 5, generator=torch.Generator().manual_seed(1))
    t = 0.02 * torch.arange(4, dtype=torch.float32)
    lo, hi = shard_rows(37, world, rank)
    sol = odeint(model, y0[lo:hi], t, method="rk4")
    grads = torch.autograd.grad(sol[-1][:, :3].sum(), list(model.parameters()))
    flat = torch.cat([g.reshape(-1) for g in grads])
    dist.all_reduce(flat)                     # the one collective of the path
    if rank == 0:
        q.put(flat.tolist())              # plain floats: a tensor would travel as an fd the exiting worker may close
    dist.barrier()
    dist.destroy_process_group()


def test_shard_rows_cover_and_balance():
    for n in (1, 2, 37, 1 << 20):
        for w in (1, 2, 4, 8):
            spans = [shard_rows(n, w, r) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def test_two_rank_allreduce_matches_single_rank_gradient():
    sys.path.insert(0, ROOT)
    from oracle import nlbac_ref
    from oracle.torchdiffeq_restated import odeint
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = torch.tensor(q.get(timeout=240))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    torch.manual_seed(0)
    model = nlbac_ref.AffineNODE(3, 3, 6)
    y0 = torch.randn(37, 5, generator=torch.Generator().manual_seed(1))
    sol = odeint(model, y0, 0.02 * torch.arange(4, dtype=torch.float32), method="rk4")
    want = torch.cat([g.reshape(-1) for g in torch.autograd.grad(sol[-1][:, :3].sum(), list(model.parameters()))])
    assert ((got - want).abs().max() / want.abs().max()) < 1e-5
