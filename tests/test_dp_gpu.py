"""Data parallelism on the CUDA path: N ranks (one per GPU) each solve their shard of the batch with the
fused kernels, d-theta is all-reduced through the C ABI's NCCL communicator (nlbac_comm_init / nlbac_allreduce_grads)
as ONE flat buffer, and must equal the single-rank d-theta of the whole batch -- tolerance 1e-5 (fp32 sums in a
different order).  Also train_step_dp (sharded model fit) against train_step over 3 Adam steps.
Needs >= 2 GPUs (skipped on a one-GPU box; run with gpurun --gpus 2)."""
import os
import sys

import pytest
import torch
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
B = 5000


def _inputs():
    g = torch.Generator().manual_seed(1)
    y0 = torch.randn(B, 5, generator=g)
    nxt = y0[:, :3] + 0.02 * torch.randn(B, 3, generator=g)
    return y0, nxt


def _model(nb):
    torch.manual_seed(0)
    return nb.NeuralODEModel(3, 3, 6)


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    import nlbac_b200 as nb
    from nlbac_b200 import parallel
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)       # only carries the 128-byte NCCL id
    comm = parallel.Comm(world, rank)
    model = _model(nb).cuda()
    y0, nxt = _inputs()
    lo, hi = parallel.shard_rows(B, world, rank)
    t = 0.02 * torch.arange(4, dtype=torch.float32)
    sol = nb.odeint(model, y0[lo:hi].cuda(), t, method="rk4")
    grads = torch.autograd.grad(sol[-1][:, :3].pow(2).sum(), list(model.parameters()))
    flat = parallel.allreduce_grad_tensors(grads, comm).clone()
    # sharded model fit, 3 Adam steps
    opt = nb.FusedAdam(model.parameters(), lr=1e-3)
    losses = []
    for _ in range(3):
        losses.append(parallel.train_step_dp(model, y0[lo:hi, :3].cuda().contiguous(), y0[lo:hi, 3:].cuda().contiguous(),
                                             nxt[lo:hi].cuda(), opt, torch.nn.MSELoss(), B, "euler", 0.02,
                                             rows_global=B, comm=comm))
    torch.cuda.synchronize()
    if rank == 0:
        q.put((flat.cpu().tolist(), losses, torch.cat([p.detach().reshape(-1) for p in model.parameters()]).cpu().tolist()))
    dist.barrier()
    comm.close()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_n_rank_allreduced_dtheta_matches_one_rank():
    import nlbac_b200 as nb
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29700 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    flat, losses, params = q.get(timeout=600)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    rel = lambda a, b: ((a - b).abs().max() / b.abs().max()).item()  # noqa: E731
    model = _model(nb).cuda()
    y0, nxt = _inputs()
    t = 0.02 * torch.arange(4, dtype=torch.float32)
    sol = nb.odeint(model, y0.cuda(), t, method="rk4")
    want = torch.cat([g.reshape(-1) for g in torch.autograd.grad(sol[-1][:, :3].pow(2).sum(), list(model.parameters()))])
    assert rel(torch.tensor(flat), want.cpu()) <= 1e-5, rel(torch.tensor(flat), want.cpu())
    opt = nb.FusedAdam(model.parameters(), lr=1e-3)
    for i in range(3):
        l1 = nb.train_step(model, y0[:, :3].cuda().contiguous(), y0[:, 3:].cuda().contiguous(), nxt.cuda(), opt,
                           torch.nn.MSELoss(), B, "euler", 0.02)
        assert abs(l1 - losses[i]) <= 1e-5 * abs(l1)
    p1 = torch.cat([p.detach().reshape(-1) for p in model.parameters()]).cpu()
    assert rel(torch.tensor(params), p1) <= 1e-5, rel(torch.tensor(params), p1)
