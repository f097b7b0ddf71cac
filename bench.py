#!/usr/bin/env python
"""bench.py — neural-ODE trajectory-steps/sec (forward + backward) on B200.

Workload (BASELINE.json configs[4], the configuration the metric is quoted on): M = 2^20 Unicycle-shaped
trajectories PER GPU x 20 rk4 (3/8-rule) steps, forward + backward with cotangent = ones on the final
state, parameter gradients produced (and all-reduced over NCCL when N > 1).  One "step" = one such
forward+backward pass.  One trajectory-step = one solver step of one trajectory.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--log2m 20] [--env U|P]

N > 1 is launched by torchrun (one rank per GPU).  Rank 0 prints ONE JSON line.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

FLOP_EVAL_FWD = {"U": 103012, "P": 106024}      # SURVEY.md App. C (1 MAC = 2 FLOP)
N_STEPS = 20
DT = 0.02


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--log2m", type=int, default=20, help="log2 of trajectories per GPU")
    ap.add_argument("--env", default="U", choices=["U", "P"])
    ap.add_argument("--precision", default="fp32", choices=["fp32", "tf32", "3xtf32"],
                    help="forward tier: fp32 FMA, tcgen05 TF32 (1e-2 tier) or 3xTF32 (fp32-class); backward is FP32")
    ap.add_argument("--cpu-sample", type=int, default=1024, help="trajectories in the CPU baseline sample")
    return ap.parse_args()


def flop_per_traj_step(env):
    return 4 * 3 * FLOP_EVAL_FWD[env]            # rk4: 4 evaluations, fwd + bwd = 3x fwd


def make_oracle_model(env):
    from oracle import nlbac_ref
    torch.manual_seed(0)
    return nlbac_ref.AffineNODE(3, 3, 6) if env == "U" else nlbac_ref.AffineNODE(6, 6, 12)


def synth_y0(env, M, seed, device):
    g = torch.Generator(device=device).manual_seed(seed)
    D = 5 if env == "U" else 8
    return torch.rand(M, D, generator=g, device=device) * 2 - 1


def cpu_reference_run(env, sample, steps, warmup, threads):
    """The reference's CPU path (oracle restatement of torchdiffeq + the reference NODE forward),
    PyTorch CPU fp32, on a bounded sample of the same workload.  Returns (traj-steps/s, ms per pass)."""
    from oracle.torchdiffeq_restated import odeint as oracle_odeint
    torch.set_num_threads(threads)
    model = make_oracle_model(env)
    y0 = synth_y0(env, sample, 1234, "cpu")
    t = DT * torch.arange(N_STEPS + 1, dtype=torch.float32)
    n_s = 3 if env == "U" else 6

    def one():
        model.zero_grad()
        sol = oracle_odeint(model, y0, t, method="rk4", atol=1e-7, rtol=1e-5)
        sol[-1][:, :n_s].sum().backward()

    for _ in range(warmup):
        one()
    t0 = time.perf_counter()
    for _ in range(steps):
        one()
    dt = (time.perf_counter() - t0) / steps
    return sample * N_STEPS / dt, dt * 1e3


class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.05)

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def fp32_peak_tflops(nlbac_b200):
    """Measured FP32-FMA peak of this GPU (register-resident FMA loop; best of 5), in TFLOP/s."""
    import ctypes
    L = nlbac_b200._lib.lib()
    sink = torch.zeros(4, device="cuda")
    flops = ctypes.c_double()
    st = torch.cuda.current_stream().cuda_stream
    best = 0.0
    for i in range(6):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        nlbac_b200._lib.check(L.nlbac_fp32_peak_probe(4096, sink.data_ptr(), ctypes.byref(flops), st), "peak_probe")
        e1.record()
        torch.cuda.synchronize()
        if i:
            best = max(best, flops.value / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    return best


def run_ours(args):
    import nlbac_b200
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    assert world == args.gpus, "launch with torchrun --nproc-per-node %d" % args.gpus
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    env, M = args.env, 1 << args.log2m
    n_s = 3 if env == "U" else 6
    ref = make_oracle_model(env)
    model = nlbac_b200.NeuralODEModel(3, 3, 6) if env == "U" else nlbac_b200.NeuralODEModel(6, 6, 12)
    model.load_state_dict(ref.state_dict())
    model.to(dev)
    params = list(model.parameters())
    t = [DT * i for i in range(N_STEPS + 1)]
    y0 = synth_y0(env, M, 1234 + rank, dev)
    y0_host = y0.cpu().pin_memory()
    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731

    def fwd_bwd(y, timers=None):
        if timers is not None:
            timers[0].record()
        sol = nlbac_b200.odeint(model, y, t, method="rk4", atol=1e-7, rtol=1e-5, precision=args.precision)
        loss = sol[-1][:, :n_s].sum()
        if timers is not None:
            timers[1].record()
        grads = torch.autograd.grad(loss, params)
        flat = grads[0]._base if grads[0]._base is not None else torch.cat([g.reshape(-1) for g in grads])
        if timers is not None:
            timers[2].record()
        if world > 1:
            dist.all_reduce(flat)
        return loss, flat

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        fwd_bwd(y0)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    timers = [[ev(), ev(), ev()] for _ in range(args.steps)]
    e0, e1 = ev(), ev()
    e0.record()
    for i in range(args.steps):
        fwd_bwd(y0, timers[i])
    e1.record()
    barrier()
    sampler.stop_flag = True
    ms_total = e0.elapsed_time(e1)
    fwd_ms = sum(tm[0].elapsed_time(tm[1]) for tm in timers) / args.steps
    bwd_ms = sum(tm[1].elapsed_time(tm[2]) for tm in timers) / args.steps

    # end to end through the public API with host buffers (pinned H2D of y0, D2H of loss + d-theta)
    y_dev = torch.empty_like(y0)
    barrier()
    h0, h1 = ev(), ev()
    h0.record()
    for _ in range(args.steps):
        y_dev.copy_(y0_host, non_blocking=True)
        loss, flat = fwd_bwd(y_dev)
        loss_h, flat_h = loss.item(), flat.cpu()
    h1.record()
    barrier()
    e2e_ms = h0.elapsed_time(h1)

    tms = torch.tensor([ms_total, e2e_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    ms_total, e2e_ms = tms.tolist()
    peak32 = fp32_peak_tflops(nlbac_b200) if rank == 0 else 0.0
    if rank == 0:
        traj_steps = float(world) * M * N_STEPS
        ms_step = ms_total / args.steps
        value = traj_steps / (ms_step * 1e-3)
        fl = flop_per_traj_step(env)
        bwd_flops = M * N_STEPS * fl * 2.0 / 3.0
        achieved = bwd_flops / (bwd_ms * 1e-3) / 1e12
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(
            os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
        cores = os.cpu_count()
        cpu_val, cpu_ms = cpu_reference_run(env, args.cpu_sample, 2, 1, cores) if world == 1 else (None, None)
        line = {
            "metric": "neural-ODE trajectory-steps/sec (fwd+bwd)", "value": value, "unit": "trajectory-steps/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "config5: 2^%d %s-shaped trajectories per GPU x 20 rk4(3/8) steps, fwd+bwd, "
                                   "t=0.02*arange(21), cotangent ones on final state, d-theta all-reduced" % (
                                       args.log2m, "Unicycle" if env == "U" else "PVTOL"),
                       "trajectories_per_gpu": M, "solver_steps": N_STEPS, "method": "rk4", "precision": {"fp32": "fwd fp32 / bwd fp32", "tf32": "fwd tf32 / bwd tf32 (dW bf16 operands, fp32 accumulate)", "3xtf32": "fwd 3xtf32 / bwd fp32"}[args.precision],
                       "l2": "per-step working set (stage checkpoints + outputs, %.1f GB) exceeds the 126 MB L2; no explicit flush" % (
                           M * (N_STEPS * 4 * n_s + (N_STEPS + 1) * (n_s + 2) * 2) * 4 / 1e9)},
            "fwd_ms": fwd_ms, "bwd_ms": bwd_ms,
            "roofline": {"bound": "fp32_fma", "kernel": "node_bwd_kernel", "achieved": achieved, "peak": peak32,
                         "unit": "TFLOP/s", "frac": achieved / peak32 if peak32 else None, "traffic": None,
                         "peak_source": "measured live: register-resident FFMA loop (nlbac_fp32_peak_probe); "
                                        "MEASURED_PEAKS.json has no FP32-FMA figure",
                         "algorithmic_flops_per_launch": bwd_flops,
                         "bf16_tensor_peak_measured_tflops": peaks.get("bf16_tflops_sustained")},
            "cpu_baseline": {"value": cpu_val, "unit": "trajectory-steps/s", "cores": cores, "kind": "port",
                             "sample": "%d trajectories x 20 rk4 steps fwd+bwd, PyTorch CPU fp32, %.0f ms per pass" % (
                                 args.cpu_sample, cpu_ms)} if cpu_val else None,
            "e2e": {"value": traj_steps / (e2e_ms / args.steps * 1e-3), "unit": "trajectory-steps/s",
                    "h2d_bytes_per_step": y0_host.numel() * 4, "d2h_bytes_per_step": 4 + flat_h.numel() * 4},
            "gpu_launches": 3 * args.steps,
            "clocks": sampler.summary(),
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def run_reference(args):
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    cores = os.cpu_count()
    val, ms = cpu_reference_run(args.env, args.cpu_sample, args.steps, args.warmup, cores)
    sample = "%d trajectories x 20 rk4 steps fwd+bwd per step, PyTorch CPU fp32" % args.cpu_sample
    print(json.dumps({
        "impl": "reference", "metric": "neural-ODE trajectory-steps/sec (fwd+bwd)", "value": val,
        "unit": "trajectory-steps/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": "config5 sample: %s" % sample, "method": "rk4", "solver_steps": N_STEPS},
        "cpu_baseline": {"value": val, "unit": "trajectory-steps/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "trajectory-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
