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
DP5_T = torch.tensor([0.0, 0.3, 0.7])


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
    # adaptive dopri5 on the sharded batch: the error norms are all-reduced, every rank takes the global step sequence
    model2 = _model(nb).cuda()
    st = {}
    sol5 = nb.odeint(model2, y0[lo:hi].cuda(), DP5_T, method="dopri5", rtol=1e-5, atol=1e-7,
                     options={"_stats": st, "comm": comm, "rows_global": B})
    g5 = torch.autograd.grad(sol5[1:, :, :3].pow(2).sum(), list(model2.parameters()))
    flat5 = parallel.allreduce_grad_tensors(g5, comm).clone()
    torch.cuda.synchronize()
    if rank == 0:
        q.put((flat.cpu().tolist(), losses, torch.cat([p.detach().reshape(-1) for p in model.parameters()]).cpu().tolist(),
               (st["n_accepted"], st["n_rejected"], st["step_dt"].tolist(), sol5.cpu().tolist(), flat5.cpu().tolist())))
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
    flat, losses, params, dp5 = q.get(timeout=600)
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
    # sharded dopri5 == the single-rank solve of the whole batch (same steps; sums in a different order)
    n_acc, n_rej, dts, sol_shard, flat5 = dp5
    model2 = _model(nb).cuda()
    st = {}
    sol = nb.odeint(model2, y0.cuda(), DP5_T, method="dopri5", rtol=1e-5, atol=1e-7, options={"_stats": st})
    want5 = torch.cat([g.reshape(-1) for g in torch.autograd.grad(sol[1:, :, :3].pow(2).sum(), list(model2.parameters()))])
    assert (n_acc, n_rej) == (st["n_accepted"], st["n_rejected"]) and n_acc >= 3
    assert rel(torch.tensor(dts), torch.tensor(st["step_dt"].tolist())) <= 1e-5
    lo, hi = nb.parallel.shard_rows(B, 2, 0)
    assert rel(torch.tensor(sol_shard), sol[:, lo:hi].detach().cpu()) <= 1e-5
    assert rel(torch.tensor(flat5), want5.cpu()) <= 5e-5, rel(torch.tensor(flat5), want5.cpu())
