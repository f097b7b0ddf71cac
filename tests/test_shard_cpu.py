"""World-size-2 gloo tests of the PACKAGE's data-parallel host logic (nlbac_b200/parallel.py): the row
sharding rule, the one-flat-buffer gradient all-reduce, and the differentiable all-reduce of the [m] constraint sums.
On CPU a small torch network stands in for the fused kernels (they need a GPU); what is under test is that a
mean-type loss scaled by shard_weight plus a loss that is NON-LINEAR in all-reduced row sums (the shape of the
CBF/CLF augmented Lagrangian, U sac_cbf_clf.py:493-528) reproduces the single-rank gradient after allreduce_grads.
The same identity on the CUDA path with the C ABI's NCCL communicator: tests/test_dp_gpu.py."""
import os
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
N_ROWS = 37


def _net():
    torch.manual_seed(0)
    return torch.nn.Sequential(torch.nn.Linear(5, 16), torch.nn.ReLU(), torch.nn.Linear(16, 3))


def _loss(net, x, rows_global, reduce_sum, weight):
    """mean-type term + a term non-linear in batch sums of where(> 0)-filtered rows."""
    out = net(x)
    mean_term = out.pow(2).mean() * weight
    rows = torch.where(out > 0, out, torch.zeros_like(out)).sum(0) / rows_global      # [m] local sums / global batch
    s = reduce_sum(rows)
    ratio = float(s[:-1].mean().abs() / s[-1].abs().clamp_min(1e-6))                  # detached, like the reference
    return mean_term + (0.3 * s + 2.0 * s * s).sum() + ratio * s[-1]


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    from nlbac_b200 import parallel
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    net = _net()
    x = torch.randn(N_ROWS, 5, generator=torch.Generator().manual_seed(1))
    lo, hi = parallel.shard_rows(N_ROWS, world, rank)
    loss = _loss(net, x[lo:hi], N_ROWS, parallel.all_reduce_sum, parallel.shard_weight(hi - lo, N_ROWS))
    loss.backward()
    flat = parallel.allreduce_grads(net)                       # the one collective per optimiser
    got = torch.cat([p.grad.reshape(-1) for p in net.parameters()])
    assert torch.equal(flat, got)
    if rank == 0:
        q.put(got.tolist())               # plain floats: a tensor would travel as an fd the exiting worker may close
    dist.barrier()
    dist.destroy_process_group()


def test_shard_rows_cover_and_balance():
    sys.path.insert(0, ROOT)
    from nlbac_b200.parallel import shard_rows
    for n in (1, 2, 37, 1 << 20):
        for w in (1, 2, 4, 8):
            spans = [shard_rows(n, w, r) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def test_two_rank_allreduce_matches_single_rank_gradient():
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
    net = _net()
    x = torch.randn(N_ROWS, 5, generator=torch.Generator().manual_seed(1))
    _loss(net, x, N_ROWS, lambda r: r, 1.0).backward()
    want = torch.cat([p.grad.reshape(-1) for p in net.parameters()])
    assert ((got - want).abs().max() / want.abs().max()) < 1e-5
