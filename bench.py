#!/usr/bin/env python
"""bench.py — neural-ODE trajectory-steps/sec (forward + backward) on B200.

Headline tier: --precision tf32 (tensor cores, gradients within 1e-2 of fp64 — measured 3.4e-3);
the exact FP32 tier (1e-5) is measured in the same run and reported under "fp32_tier".

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


def ncu_traffic(precision):
    """dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of the dominant kernel at the bench size, read from
    the `ncu --set full` summary committed under profiles/ (None when no capture exists for that tier)."""
    import re
    f, kern = {"tf32": ("r1_ncu_bench_tc_tf32.txt", "node_tc_bwd_kernel"),
               "fp32": (None, None), "3xtf32": (None, None)}[precision]   # no ncu --set full capture of the other tiers
    path = os.path.join(ROOT, "profiles", f) if f else None
    if not path or not os.path.exists(path):
        return None, None
    unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    tot, inside = 0.0, False
    for line in open(path):
        if line.startswith("kernel:"):
            if inside:
                break
            inside = kern in line
        elif inside:
            m = re.match(r"dram__bytes_(read|write)\.sum\s+(\w+)\s+([0-9.]+)", line)
            if m:
                tot += float(m.group(3)) * unit[m.group(2)]
    return (tot or None), "profiles/" + f


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--log2m", type=int, default=20, help="log2 of trajectories per GPU")
    ap.add_argument("--env", default="U", choices=["U", "P"])
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: 2^log2m trajectories PER GPU; strong: 2^log2m-total trajectories sharded M/world per GPU")
    ap.add_argument("--log2m-total", type=int, default=23, help="log2 of the total trajectories under --scaling strong")
    ap.add_argument("--precision", default="tf32", choices=["fp32", "tf32", "3xtf32"],
                    help="tf32: tcgen05 kind::tf32 forward + reverse sweep, dW in bf16 (the 1e-2 tier of north_star); "
                         "fp32: exact FP32 FMA path (1e-5 tier); 3xtf32: tcgen05 with hi/lo split operands end to end")
    ap.add_argument("--cpu-sample", type=int, default=65536,
                    help="trajectories in the CPU baseline sample (2^20 x 25 passes does not fit a bench run: ~40 s per "
                         "pass on 16 cores; the CPU rate still grows with the sample -- 1.5e5 at 2^10, 4.8e5 at 2^14 --, "
                         "cpu_baseline.flatness reports a second point)")
    ap.add_argument("--sweep-only", action="store_true",
                    help="time the headline tier only (no FP32 / 3xTF32 tiers, no baselines, no peak probe): for the "
                         "1M-64M sweep of BASELINE config 5")
    ap.add_argument("--suite", default="headline", choices=["headline", "configs"],
                    help="configs: latencies of the reference's real small-batch configurations 1-3 (SURVEY.md §8d) "
                         "next to the same code run eagerly by PyTorch on the GPU and on the host cores")
    return ap.parse_args()


def flop_per_traj_step(env):
    return 4 * 3 * FLOP_EVAL_FWD[env]            # rk4: 4 evaluations, fwd + bwd = 3x fwd


REF_VARIANT = {"U": "NLBAC_Unicycle_RL_training/Unicycle_RL_training", "P": "NLBAC_pvtol_RL_training/Pvtol_RL_training"}


def make_reference_model(env, device="cpu"):
    """The NODE of the reference arm under torch.manual_seed(0), and what it is: the reference's OWN
    ``sac_cbf_clf/model.py`` class imported by path (kind "reference": only where /root/reference exists, i.e. in the
    build container -- its module-global ``device`` is patched, U model.py:11, and ``torchdiffeq`` is the restated
    solver, not installable anywhere we run), else the oracle's restatement of that class (kind "port")."""
    import importlib.util
    import types
    from oracle import nlbac_ref, torchdiffeq_restated
    path = os.path.join("/root/reference", REF_VARIANT[env], "sac_cbf_clf", "model.py")
    dims = (3, 3, 6) if env == "U" else (6, 6, 12)
    if os.path.exists(path):
        stub = types.ModuleType("torchdiffeq")
        stub.odeint = torchdiffeq_restated.odeint
        sys.modules.setdefault("torchdiffeq", stub)
        spec = importlib.util.spec_from_file_location("nlbac_reference_model_" + env, path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        mod.device = torch.device(device)
        torch.manual_seed(0)
        return mod.NeuralODEModel(*dims).to(device), "reference"
    torch.manual_seed(0)
    return nlbac_ref.AffineNODE(*dims).to(device), "port"


def synth_y0(env, M, seed, device):
    g = torch.Generator(device=device).manual_seed(seed)
    D = 5 if env == "U" else 8
    return torch.rand(M, D, generator=g, device=device) * 2 - 1


def reference_run(env, sample, steps, warmup, threads=None, device="cpu"):
    """The reference's path (its NODE class, see make_reference_model, + the restated torchdiffeq rk4), eager PyTorch
    fp32 (TF32 off) on a bounded sample of the same workload; on the host cores (``threads``) or, with
    device="cuda", run eagerly on the GPU -- SURVEY.md 8(d)'s like-for-like bar.
    Returns (traj-steps/s, ms per pass, kind)."""
    from oracle.torchdiffeq_restated import odeint as oracle_odeint
    if threads:
        torch.set_num_threads(threads)
    torch.backends.cuda.matmul.allow_tf32 = False
    model, kind = make_reference_model(env, device)
    y0 = synth_y0(env, sample, 1234, device)
    t = (DT * torch.arange(N_STEPS + 1, dtype=torch.float32)).to(device)
    n_s = 3 if env == "U" else 6
    sync = torch.cuda.synchronize if device != "cpu" else (lambda: None)

    def one():
        model.zero_grad()
        sol = oracle_odeint(model, y0, t, method="rk4", atol=1e-7, rtol=1e-5)
        sol[-1][:, :n_s].sum().backward()

    for _ in range(warmup):
        one()
    sync()
    t0 = time.perf_counter()
    for _ in range(steps):
        one()
    sync()
    dt = (time.perf_counter() - t0) / steps
    return sample * N_STEPS / dt, dt * 1e3, kind


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


def kernel_timing(nlbac_b200, enable):
    nlbac_b200._lib.check(nlbac_b200._lib.lib().nlbac_kernel_timing(1 if enable else 0), "kernel_timing")


def kernel_timing_read(nlbac_b200):
    import ctypes
    ms, n = (ctypes.c_double * 4)(), (ctypes.c_int * 4)()
    nlbac_b200._lib.check(nlbac_b200._lib.lib().nlbac_kernel_timing_read(ms, n), "kernel_timing_read")
    return list(ms), list(n)


def image_bytes_per_tile_eval(nlbac_b200, model, D, method="rk4", prec=2):
    """bf16 image bytes pass A writes (and pass B reads) per 128-trajectory tile per stage evaluation."""
    import ctypes
    from nlbac_b200.odeint import describe, METHODS
    L = nlbac_b200._lib.lib()
    desc, _ = describe(model, D)
    out = []
    for B in (128, 256):
        nb = ctypes.c_size_t()
        nlbac_b200._lib.check(L.nlbac_node_tc_bwd_workspace_bytes(ctypes.byref(desc), B, 2, 2, METHODS[method], prec, 1,
                                                                  ctypes.byref(nb)), "tc_ws")
        out.append(nb.value)
    return (out[1] - out[0]) / 4.0           # one rk4 step = 4 stage evaluations


def run_ours(args):
    import nlbac_b200
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    assert world == args.gpus, "launch with torchrun --nproc-per-node %d" % args.gpus
    # stdout carries exactly ONE JSON line: anything libraries print on fd 1 (NCCL's version banner) goes to stderr
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    env = args.env
    if args.scaling == "strong":
        lo_, hi_ = nlbac_b200.parallel.shard_rows(1 << args.log2m_total, world, rank)
        M = hi_ - lo_
    else:
        M = 1 << args.log2m
    comm = nlbac_b200.parallel.Comm(world, rank) if world > 1 else None      # the C ABI's NCCL communicator
    n_s = 3 if env == "U" else 6
    torch.manual_seed(0)     # same nn.Linear construction order as the reference class -> same default init
    model = nlbac_b200.NeuralODEModel(3, 3, 6) if env == "U" else nlbac_b200.NeuralODEModel(6, 6, 12)
    model.to(dev)
    params = list(model.parameters())
    t = [DT * i for i in range(N_STEPS + 1)]
    y0 = synth_y0(env, M, 1234 + rank, dev)
    y0_host = y0.cpu().pin_memory()
    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731

    def fwd_bwd(y, timers=None, precision=args.precision):
        if timers is not None:
            timers[0].record()
        sol = nlbac_b200.odeint(model, y, t, method="rk4", atol=1e-7, rtol=1e-5, precision=precision)
        loss = sol[-1][:, :n_s].sum()
        if timers is not None:
            timers[1].record()
        grads = torch.autograd.grad(loss, params)
        if timers is not None:
            timers[2].record()
        # ONE flat d-theta buffer (the backward kernel writes all gradients into one allocation), ONE all-reduce
        flat = nlbac_b200.parallel.allreduce_grad_tensors(grads, comm)
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
    kernel_timing(nlbac_b200, True)
    e0, e1 = ev(), ev()
    e0.record()
    for i in range(args.steps):
        fwd_bwd(y0, timers[i])
    e1.record()
    barrier()
    sampler.stop_flag = True
    k_ms, k_n = kernel_timing_read(nlbac_b200)      # summed over the K timed steps, per kernel slot
    kernel_timing(nlbac_b200, False)
    ms_total = e0.elapsed_time(e1)
    fwd_ms = sum(tm[0].elapsed_time(tm[1]) for tm in timers) / args.steps
    bwd_ms = sum(tm[1].elapsed_time(tm[2]) for tm in timers) / args.steps

    # end to end through the public API with host buffers (pinned H2D of y0, D2H of loss + d-theta)
    # every step: H2D of that step's y0 from pinned memory, D2H of its loss and d-theta into pinned memory (asynchronous
    # copies on the compute stream, inside the timed region; one synchronisation at the end, like a training loop that
    # logs from the previous step -- a blocking .item() per step would put host jitter into the figure)
    y_dev = torch.empty_like(y0)
    loss_h = torch.empty(1, dtype=torch.float32).pin_memory()
    flat_h = torch.empty(sum(p.numel() for p in params), dtype=torch.float32).pin_memory()
    barrier()
    h0, h1 = ev(), ev()
    h0.record()
    for _ in range(args.steps):
        y_dev.copy_(y0_host, non_blocking=True)
        loss, flat = fwd_bwd(y_dev)
        loss_h.copy_(loss.detach().reshape(1), non_blocking=True)
        flat_h.copy_(flat, non_blocking=True)
    h1.record()
    barrier()
    assert torch.isfinite(loss_h).all() and torch.isfinite(flat_h).all()
    e2e_ms = h0.elapsed_time(h1)

    # the exact FP32 tier on the same workload (1 warm-up + 2 timed passes), reported beside the headline
    fp32_ms = 0.0
    if args.precision != "fp32" and not args.sweep_only:
        fwd_bwd(y0, precision="fp32")
        barrier()
        f0, f1 = ev(), ev()
        f0.record()
        for _ in range(2):
            fwd_bwd(y0, precision="fp32")
        f1.record()
        barrier()
        fp32_ms = f0.elapsed_time(f1) / 2

    # the 3xTF32 tier (forward, reverse sweep and dW all on tcgen05 with hi/lo split operands), same workload
    x3_ms = 0.0
    if args.precision != "3xtf32" and not args.sweep_only:
        fwd_bwd(y0, precision="3xtf32")
        barrier()
        f0, f1 = ev(), ev()
        f0.record()
        for _ in range(2):
            fwd_bwd(y0, precision="3xtf32")
        f1.record()
        barrier()
        x3_ms = f0.elapsed_time(f1) / 2

    tms = torch.tensor([ms_total, e2e_ms, fp32_ms, x3_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    ms_total, e2e_ms, fp32_ms, x3_ms = tms.tolist()
    peak32 = fp32_peak_tflops(nlbac_b200) if rank == 0 else 0.0
    if rank == 0:
        traj_steps = float((1 << args.log2m_total) if args.scaling == "strong" else world * M) * N_STEPS
        ms_step = ms_total / args.steps
        value = traj_steps / (ms_step * 1e-3)
        K = args.steps
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(
            os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
        evals = M * N_STEPS * 4.0                               # stage evaluations per pass (rk4)
        sweep_ms, sweep_n = k_ms[1] / K, k_n[1] // K            # reverse sweep: the dominant kernel in both tiers
        if args.precision in ("tf32", "3xtf32"):
            # algorithmic work of pass A = the dX products (1 x forward FLOPs); it also re-computes the forward
            # hidden layers from the stage checkpoints (another 1 x), reported as "executed"
            sweep_flops, executed = evals * FLOP_EVAL_FWD[env], evals * 2 * FLOP_EVAL_FWD[env]
            bf16 = peaks.get("bf16_tflops_sustained") or 1403.6
            peak, peak_src = bf16 / 2.0, (
                "MEASURED_PEAKS.json bf16_tflops_sustained / 2 (kind::tf32 issues at half the bf16 rate; sustained "
                "figure because the kernel is timed inside a long step)" if peaks else
                "B200_PROFILING.md fallback: sustained bf16 1403.6 / 2")
            if args.precision == "3xtf32":      # three MMAs per algorithmic product (TF32 peak / 3)
                peak, peak_src = peak / 3.0, peak_src + "; / 3 for the 3xTF32 split (hi*hi + lo*hi + hi*lo)"
            roof = {"bound": "tensor", "kernel": "node_tc_bwd_kernel (pass A: tcgen05 %s recompute + dX sweep)" % (
                "3xTF32" if args.precision == "3xtf32" else "TF32")}
            img = image_bytes_per_tile_eval(nlbac_b200, model, 5 if env == "U" else 8, prec=2 if args.precision == "tf32" else 3)
            dw_bytes = (M / 128.0) * N_STEPS * 4 * img
            hbm = peaks.get("hbm_gbs") or 6453.1
            dw_gbs = dw_bytes / (k_ms[2] / K * 1e-3) / 1e9 if k_ms[2] else None
            secondary = {"kernel": "node_tc_dw_kernel (pass B: bf16 dW = d^T a over the HBM images)", "bound": "hbm",
                         "achieved": dw_gbs, "peak": hbm, "unit": "GB/s", "frac": dw_gbs / hbm if dw_gbs else None,
                         "image_bytes_per_pass": dw_bytes,
                         "algorithmic_bytes_per_pass": M * (20.0 + (N_STEPS + 1) * 20.0 + 2 * N_STEPS * 4 * 12.0 + 2 * 20.0)
                         if env == "U" else None, "ms_per_pass": k_ms[2] / K,
                         "launches_per_pass": k_n[2] // K}
        else:
            # algorithmic work = dX + dW (2 x forward FLOPs); the kernel also re-computes the forward (3 x executed)
            sweep_flops, executed = evals * 2 * FLOP_EVAL_FWD[env], evals * 3 * FLOP_EVAL_FWD[env]
            peak, peak_src = peak32, ("measured live: register-resident FFMA loop (nlbac_fp32_peak_probe); "
                                      "MEASURED_PEAKS.json has no FP32-FMA figure")
            roof = {"bound": "fp32_fma", "kernel": "node_bwd_kernel"}
            secondary = None
        achieved = sweep_flops / (sweep_ms * 1e-3) / 1e12 if sweep_ms else None
        roof.update({"achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                     "frac": achieved / peak if achieved and peak else None, "traffic": ncu_traffic(args.precision)[0], "traffic_source": ncu_traffic(args.precision)[1],
                     "peak_source": peak_src, "algorithmic_flops_per_pass": sweep_flops,
                     "executed_tflops": executed / (sweep_ms * 1e-3) / 1e12 if sweep_ms else None,
                     "ms_per_pass": sweep_ms, "launches_per_pass": sweep_n,
                     "share_of_step": sweep_ms / ms_step if ms_step else None})
        # SURVEY.md 8(d)'s whole-path figure: trajectory-steps/s x algorithmic FLOPs per trajectory-step (fwd + bwd =
        # 3 x forward, recompute not counted) against the same peak
        whole = value / world * flop_per_traj_step(env) / 1e12
        roof["whole_step"] = {"achieved": whole, "unit": "TFLOP/s", "frac": whole / peak if peak else None}
        cores = os.cpu_count()
        cpu_base = eager_gpu = None
        if world == 1 and not args.sweep_only:
            del y_dev
            torch.cuda.empty_cache()
            cs = args.cpu_sample
            cv, cms, kind = reference_run(env, cs, 2, 1, cores)
            small = max(64, cs // 4)
            fv, fms, _ = reference_run(env, small, 2, 1, cores)
            ov, oms, _ = reference_run(env, min(cs, 2048), 1, 1, 1)
            cpu_base = {"value": cv, "unit": "trajectory-steps/s", "cores": cores, "kind": kind,
                        "sample": "%d trajectories x 20 rk4 steps fwd+bwd (the reference's NeuralODEModel + restated "
                                  "torchdiffeq rk4, PyTorch CPU fp32), %.0f ms per pass, 1 warm-up + 2 timed" % (cs, cms),
                        "flatness": {"sample": small, "value": fv, "ms_per_pass": fms},
                        "one_thread": {"sample": min(cs, 2048), "value": ov, "ms_per_pass": oms, "cores": 1}}
            gs = min(M, 1 << 16)      # eager autograd keeps ~300 KB of activations per trajectory
            gv, gms, gkind = reference_run(env, gs, 2, 1, None, device="cuda")
            eager_gpu = {"value": gv, "unit": "trajectory-steps/s", "kind": gkind, "ms_per_pass": gms,
                         "sample": "%d trajectories x 20 rk4 steps fwd+bwd: the same reference classes + restated "
                                   "torchdiffeq run eagerly by PyTorch on this B200 (fp32, TF32 off)" % gs}
        if args.precision == "tf32" and not args.sweep_only:
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            import umma_peak
            pk = umma_peak.measure(1.5, modes=(1, 3))
            tf32_peak = {"pipe": pk["tf32_ts_n256"], "kernel_tile_n112": pk["tf32_ts_n112"],
                         "how": "nlbac_umma_peak_probe: one thread per SM issues tcgen05.mma kind::tf32 back to back, "
                                "A in tensor memory, B in shared memory, M=128, N=256 (pipe) / N=112 (the NODE tile); "
                                "burst = best ~10 ms launch, sustained = 1.5 s of launches"}
            roof["peak_measured_tf32"] = tf32_peak
            mp = tf32_peak["pipe"]["sustained_tflops"]
            roof["frac_of_measured_tf32"] = achieved / mp if achieved and mp else None
            roof["whole_step"]["frac_of_measured_tf32"] = whole / mp if mp else None
        line = {
            "metric": "neural-ODE trajectory-steps/sec (fwd+bwd)", "value": value, "unit": "trajectory-steps/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
            "dtype": {"fp32": "f32", "tf32": "tf32", "3xtf32": "f32 (3xtf32 fwd)"}[args.precision], "data": "synthetic",
            "config": {"workload": "config5: %s %s-shaped trajectories x 20 rk4(3/8) steps, fwd+bwd, "
                                   "t=0.02*arange(21), cotangent ones on final state, d-theta all-reduced" % (
                                       ("2^%d in total, sharded M/world per GPU" % args.log2m_total)
                                       if args.scaling == "strong" else ("2^%d per GPU" % args.log2m),
                                       "Unicycle" if env == "U" else "PVTOL"),
                       "trajectories_per_gpu": M, "solver_steps": N_STEPS, "method": "rk4", "precision": {"fp32": "fwd fp32 / bwd fp32", "tf32": "fwd tf32 / bwd tf32 (dW bf16 operands, fp32 accumulate)", "3xtf32": "fwd 3xtf32 / bwd 3xtf32 (dW from bf16 hi/lo image pairs, fp32 accumulate, drained per block)"}[args.precision],
                       "l2": "per-step working set (stage checkpoints + outputs, %.1f GB) exceeds the 126 MB L2; no explicit flush" % (
                           M * (N_STEPS * 4 * n_s + (N_STEPS + 1) * (n_s + 2) * 2) * 4 / 1e9)},
            "fwd_ms": fwd_ms, "bwd_ms": bwd_ms,
            "kernel_ms_per_step": {"fwd_solve": k_ms[0] / K, "reverse_sweep": k_ms[1] / K, "dw_gemm": k_ms[2] / K,
                                   "grad_fold": k_ms[3] / K},
            "roofline": roof, "roofline_secondary": secondary,
            "fp32_tier": {"value": traj_steps / (fp32_ms * 1e-3), "ms_per_step": fp32_ms,
                          "note": "exact FP32 FMA path (1e-5 tier), same workload, 2 timed passes"} if fp32_ms else None,
            "tier_3xtf32": {"value": traj_steps / (x3_ms * 1e-3), "ms_per_step": x3_ms,
                            "note": "3xTF32 tier: forward, reverse sweep and dW on tcgen05 with hi/lo split operands "
                                    "(dL/dy0 2e-5, d-theta 2e-5 per tensor at B <= 8192, 5e-5 flat beyond), same workload, "
                                    "2 timed passes"} if x3_ms else None,
            "cpu_baseline": cpu_base, "eager_gpu_baseline": eager_gpu,
            "e2e": {"value": traj_steps / (e2e_ms / args.steps * 1e-3), "unit": "trajectory-steps/s",
                    "h2d_bytes_per_step": y0_host.numel() * 4, "d2h_bytes_per_step": 4 + flat_h.numel() * 4},
            "gpu_launches": int(sum(k_n)),
            "clocks": sampler.summary(),
        }
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if comm is not None:
        comm.close()
    if world > 1:
        dist.destroy_process_group()


def run_reference(args):
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    cores = os.cpu_count()
    val, ms, kind = reference_run(args.env, args.cpu_sample, args.steps, args.warmup, cores)
    sample = ("%d trajectories x 20 rk4 steps fwd+bwd per step (config5 is 2^20 per GPU, ~40 s per pass on 16 cores: "
              "the arm times a bounded sample), PyTorch CPU fp32, all %d host threads" % (args.cpu_sample, cores))
    print(json.dumps({
        "impl": "reference", "metric": "neural-ODE trajectory-steps/sec (fwd+bwd)", "value": val,
        "unit": "trajectory-steps/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": "config5 sample: %s" % sample, "method": "rk4", "solver_steps": N_STEPS},
        "cpu_baseline": {"value": val, "unit": "trajectory-steps/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": val, "unit": "trajectory-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def run_configs(args):
    """BASELINE.json configs[0..2] as the reference runs them (B = 128 / 256, one solver call per update): median
    latency of forward + backward through the public API, in microseconds, for this library, for the reference's
    model classes + restated torchdiffeq run eagerly by PyTorch on the same GPU, and for the same on the host cores."""
    import statistics
    import nlbac_b200
    from oracle import nlbac_ref
    from oracle.torchdiffeq_restated import odeint as oracle_odeint
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)

    def lat_gpu(fn, iters=50, warm=10):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(iters):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        return statistics.median(ts)

    def lat_cpu(fn, iters=7, warm=2):
        for _ in range(warm):
            fn()
        ts = []
        for _ in range(iters):
            t0 = time.perf_counter()
            fn()
            ts.append((time.perf_counter() - t0) * 1e6)
        return statistics.median(ts)

    def node_case(name, mk_ref, mk_ours, B, D, n_s, method, steps=1, options=None, graphed=False):
        torch.manual_seed(0)
        ref = mk_ref()
        ours = mk_ours()
        ours.load_state_dict(ref.state_dict())
        ours.to(dev)
        ref_gpu = mk_ref()
        ref_gpu.load_state_dict(ref.state_dict())
        ref_gpu.to(dev)
        y_cpu = torch.rand(B, D) * 2 - 1
        y_gpu = y_cpu.to(dev)
        t = [DT * i for i in range(steps + 1)]
        tt = torch.tensor(t)

        def ours_fn():
            sol = nlbac_b200.odeint(ours, y_gpu, t, method=method, atol=1e-7, rtol=1e-5,
                                    options=dict(options) if options else None)
            return torch.autograd.grad(sol[-1][:, :n_s].sum(), list(ours.parameters()))

        if graphed:      # the whole forward + backward recorded once; a call is one cudaGraphLaunch
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(3):
                    ours_fn()
            torch.cuda.current_stream().wait_stream(side)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                keep = ours_fn()                                              # noqa: F841 (outputs live in the graph pool)
            eager_ours, ours_fn = ours_fn, graph.replay                       # noqa: F841
            if options and "_stats" in options:
                graph.replay()
                info = options["_stats"]["read"]()
                assert info["done"] == 1 and info["err"] == 0, info

        def eager_fn(model, y, tv):
            def f():
                sol = oracle_odeint(model, y, tv, method=method, atol=1e-7, rtol=1e-5)
                torch.autograd.grad(sol[-1][:, :n_s].sum(), list(model.parameters()))
            return f

        o = lat_gpu(ours_fn)
        e = lat_gpu(eager_fn(ref_gpu, y_gpu, tt.to(dev)), iters=20, warm=3)
        c = lat_cpu(eager_fn(ref, y_cpu, tt))
        return {"case": name, "batch": B, "method": method, "solver_steps": steps, "ours_us": o,
                "torch_eager_gpu_us": e, "torch_cpu_us": c, "speedup_vs_eager_gpu": e / o, "speedup_vs_cpu": c / o}

    U = (lambda: nlbac_ref.AffineNODE(3, 3, 6), lambda: nlbac_b200.NeuralODEModel(3, 3, 6))
    P = (lambda: nlbac_ref.AffineNODE(6, 6, 12), lambda: nlbac_b200.NeuralODEModel(6, 6, 12))
    Cc = (lambda: nlbac_ref.BlackBoxNODE(12, 10), lambda: nlbac_b200.NeuralODEModel(12, 10))
    rows = [node_case("config1 Unicycle predict", *U, 128, 5, 3, "euler"),
            node_case("config1 Unicycle predict", *U, 128, 5, 3, "rk4"),
            node_case("config1 Unicycle NODE batch", *U, 32768, 5, 3, "euler"),
            node_case("config2 SimulatedCars predict", *Cc, 256, 12, 10, "euler"),
            node_case("config3 PVTOL predict", *P, 256, 8, 6, "dopri5"),
            node_case("config3 PVTOL predict (host-side step controller)", *P, 256, 8, 6, "dopri5",
                      options={"controller": "host"}),
            node_case("config3 PVTOL predict (one CUDA-graph replay)", *P, 256, 8, 6, "dopri5",
                      options={"max_attempts": 4, "_stats": {}}, graphed=True),
            node_case("config3 PVTOL predict", *P, 256, 8, 6, "rk4", 20)]

    def mlp_case(name, mk_ref, mk_ours, B, n_in):
        torch.manual_seed(0)
        ref = mk_ref()
        ours = mk_ours()
        ours.load_state_dict(ref.state_dict())
        ours.to(dev)
        ref_gpu = mk_ref()
        ref_gpu.load_state_dict(ref.state_dict())
        ref_gpu.to(dev)
        xs_cpu = [torch.randn(B, n) for n in n_in]
        xs_gpu = [x.to(dev) for x in xs_cpu]

        def run(model, xs):
            def f():
                out = model(*xs)
                out = out if isinstance(out, tuple) else (out,)
                torch.autograd.grad(sum(o.sum() for o in out), list(model.parameters()))
            return f
        o, e, c = lat_gpu(run(ours, xs_gpu)), lat_gpu(run(ref_gpu, xs_gpu), iters=20, warm=3), lat_cpu(run(ref, xs_cpu))
        return {"case": name, "batch": B, "ours_us": o, "torch_eager_gpu_us": e, "torch_cpu_us": c,
                "speedup_vs_eager_gpu": e / o, "speedup_vs_cpu": c / o}

    rows += [mlp_case("config1 twin-Q fwd+bwd", lambda: nlbac_ref.QNetwork(7, 2, 256), lambda: nlbac_b200.QNetwork(7, 2, 256),
                      128, (7, 2)),
             mlp_case("config1 Lyapunov fwd+bwd", lambda: nlbac_ref.LyaNetwork(2, 256), lambda: nlbac_b200.LyaNetwork(2, 256),
                      128, (2,)),
             mlp_case("config2 twin-Q fwd+bwd", lambda: nlbac_ref.QNetwork(10, 1, 256), lambda: nlbac_b200.QNetwork(10, 1, 256),
                      256, (10, 1))]
    # one NODE model-fit step (train_step, U model.py:221-260): eager API, CUDA-graph replay, reference eager
    def train_case(N):
        torch.manual_seed(0)
        ref = nlbac_ref.AffineNODE(3, 3, 6)
        ref_gpu = nlbac_ref.AffineNODE(3, 3, 6).to(dev)
        ref_gpu.load_state_dict(ref.state_dict())
        ours = nlbac_b200.NeuralODEModel(3, 3, 6)
        ours.load_state_dict(ref.state_dict())
        ours.to(dev)
        graphed_model = nlbac_b200.NeuralODEModel(3, 3, 6)
        graphed_model.load_state_dict(ref.state_dict())
        graphed_model.to(dev)
        st, ac = torch.randn(N, 3), torch.randn(N, 2)
        nx = st + 0.02 * torch.randn(N, 3)
        std, acd, nxd = st.to(dev), ac.to(dev), nx.to(dev)
        mse = torch.nn.MSELoss()
        o_opt = nlbac_b200.FusedAdam(ours.parameters(), lr=1e-3)
        r_opt = torch.optim.Adam(ref_gpu.parameters(), lr=1e-3)
        c_opt = torch.optim.Adam(ref.parameters(), lr=1e-3)
        gstep = nlbac_b200.GraphedTrainStep(graphed_model, N, 3, 2, solver="euler", time_interval=DT, lr=1e-3)
        o = lat_gpu(lambda: nlbac_b200.train_step(ours, std, acd, nxd, o_opt, mse, N, "euler", DT))
        gq = lat_gpu(lambda: gstep(std, acd, nxd))
        e = lat_gpu(lambda: nlbac_ref.train_step(ref_gpu, std, acd, nxd, r_opt, mse, N, "euler", DT), iters=20, warm=3)
        c = lat_cpu(lambda: nlbac_ref.train_step(ref, st, ac, nx, c_opt, mse, N, "euler", DT))
        return {"case": "config1 train_step (fit + Adam)", "batch": N, "method": "euler", "ours_us": o,
                "ours_cuda_graph_us": gq, "torch_eager_gpu_us": e, "torch_cpu_us": c, "speedup_vs_eager_gpu": e / o,
                "speedup_vs_cpu": c / o, "graph_speedup_vs_eager_gpu": e / gq}
    rows += [train_case(128), train_case(32768)]

    # BASELINE config 4: Quadrotor-shaped (8 -> 64 x3 -> 6, hidden 64 assumed) horizon-10 chained rollout with the policy
    # re-sampled between steps (eps[9, 256, 2] injected), forward + backward to the first action and the NODE weights
    def rollout_case():
        from oracle import constraints as OC
        torch.manual_seed(0)
        B, H = 256, 10
        ref_node, ref_pol = nlbac_ref.BlackBoxNODE(8, 6), nlbac_ref.GaussianPolicy(6, 2, 256)
        node, pol = nlbac_b200.NeuralODEModel(8, 6), nlbac_b200.GaussianPolicy(6, 2, 256)
        node.load_state_dict(ref_node.state_dict()); pol.load_state_dict(ref_pol.state_dict())
        node, pol = node.to(dev), pol.to(dev)
        gn, gp = nlbac_ref.BlackBoxNODE(8, 6).to(dev), nlbac_ref.GaussianPolicy(6, 2, 256).to(dev)
        gn.load_state_dict(ref_node.state_dict()); gp.load_state_dict(ref_pol.state_dict())
        gp.action_scale, gp.action_bias = gp.action_scale.to(dev), gp.action_bias.to(dev)
        x0, eps, u0 = torch.randn(B, 6), torch.randn(H - 1, B, 2), torch.randn(B, 2)

        def run(fn, n, p_, x, e, u):
            def f():
                uu = u.detach().requires_grad_(True)
                out = fn(n, p_, x, uu, H, dt=DT, eps=e)
                torch.autograd.grad(out.sum(), [uu] + list(n.parameters()))
            return f
        o = lat_gpu(run(nlbac_b200.constraints.chained_rollout, node, pol, x0.to(dev), eps.to(dev), u0.to(dev)))
        e = lat_gpu(run(OC.chained_rollout, gn, gp, x0.to(dev), eps.to(dev), u0.to(dev)), iters=20, warm=3)
        c = lat_cpu(run(OC.chained_rollout, ref_node, ref_pol, x0, eps, u0))
        return {"case": "config4 Quadrotor-shaped rollout, horizon 10, policy re-sampled between steps (fwd+bwd)",
                "batch": B, "method": "euler", "solver_steps": H, "ours_us": o, "torch_eager_gpu_us": e, "torch_cpu_us": c,
                "speedup_vs_eager_gpu": e / o, "speedup_vs_cpu": c / o,
                "note": "hidden width 64 ASSUMED (safe-control-gym submodule missing from the reference)"}
    rows += [rollout_case()]

    # one full update_parameters (U sac_cbf_clf.py:181-319) at the reference's batch size: GraphedUpdate replay vs the
    # restated update run eagerly by PyTorch on the GPU and on the host (inputs: the committed golden trace's batches)
    def update_case():
        import numpy as np
        from oracle.update import UnicycleAgent
        g = torch.load(os.path.join(ROOT, "tests", "golden", "update_U.pt"), weights_only=False)
        k, B = g["consts"], g["B"]
        C = nlbac_b200.constraints
        sp = type("S", (), {"low": np.array(k["action_low"], dtype="float32"), "high": np.array(k["action_high"], dtype="float32")})
        nets = {"critic": nlbac_b200.QNetwork(7, 2, 256), "critic_target": nlbac_b200.QNetwork(7, 2, 256),
                "lya": nlbac_b200.LyaNetwork(2, 256), "lya_t": nlbac_b200.LyaNetwork(2, 256),
                "policy": nlbac_b200.GaussianPolicy(7, 2, 256, sp), "backup": nlbac_b200.GaussianPolicy(7, 2, 256, sp),
                "node": nlbac_b200.NeuralODEModel(3, 3, 6)}
        for n, key in (("critic", "critic"), ("critic_target", "critic"), ("lya", "lyapunovNet"), ("lya_t", "lyapunovNet"),
                       ("policy", "policy"), ("backup", "backup_policy"), ("node", "neural_ode_model")):
            nets[n].load_state_dict(g["init"][key])
            nets[n].to(dev)
        kw = dict(dt=k["dt"], gamma_b=k["gamma_b"], hazards_locations=k["hazards"], hazards_radius=k["hazards_radius"],
                  l_p=k["l_p"], batch_size=B)
        node, lya = nets["node"], nets["lya"]
        env = type("E", (), {"dynamics_mode": "Unicycle"})
        def make(concurrent):
            return nlbac_b200.GraphedUpdate(
                nets["critic"], nets["critic_target"], lya, nets["lya_t"], nets["policy"], nets["backup"], node,
                nlbac_b200.DynamicsModel(env), B,
                primary=lambda s_, pi, V: C.unicycle_required_matrix(node, lya, s_, pi, V, **kw),
                backup=lambda s_, pi: C.unicycle_required_matrix(node, None, s_, pi, None, **kw),
                num_constraints=8, obs_dim=7, act_dim=2, center_pos_num=2, concurrent=concurrent)
        batch = tuple(None if x is None else x.to(dev) for x in g["batches"][0])
        upd1 = make(False)
        upd1.step(batch)
        o_single = lat_gpu(lambda: upd1.graph.replay())                  # every launch on one stream: a linear graph
        upd = make(True)
        upd.step(batch)                                                  # capture
        o = lat_gpu(lambda: upd.step(batch))
        o_replay = lat_gpu(lambda: upd.graph.replay())
        # the call the training loop makes: sample the device-resident replay memory + update (one gather straight into
        # the static buffers) vs sample-then-copy
        mem = nlbac_b200.ReplayMemory(4 * B, 0)
        for b_ in g["batches"][:4]:
            cols = [x.numpy() for x in b_[:8]]
            for i in range(B):
                mem.push(*[c_[i] for c_ in cols])

        class SampleOnly:
            position = mem.position

            @staticmethod
            def sample(batch_size):
                return mem.sample(batch_size)
        o_mem = lat_gpu(lambda: upd.update_parameters(mem, B, 9, None, None, 10))
        o_mem_copy = lat_gpu(lambda: upd.update_parameters(SampleOnly, B, 9, None, None, 10))
        ag_gpu, ag_cpu = UnicycleAgent(k, B, g["init"], "cuda"), UnicycleAgent(k, B, g["init"], "cpu")
        e = lat_gpu(lambda: ag_gpu.update_parameters(batch, 9), iters=20, warm=3)
        c = lat_cpu(lambda: ag_cpu.update_parameters(g["batches"][0], 9))
        return {"case": "config1 full update_parameters (critic + Lyapunov + actor + backup actor + alpha + Polyak)",
                "batch": B, "ours_us": o, "ours_graph_replay_only_us": o_replay,
                "ours_graph_replay_single_stream_us": o_single, "ours_sample_and_update_us": o_mem,
                "ours_sample_copy_and_update_us": o_mem_copy, "torch_eager_gpu_us": e, "torch_cpu_us": c,
                "speedup_vs_eager_gpu": e / o, "speedup_vs_cpu": c / o,
                "note": "ours = load the minibatch into the static buffers + ONE CUDA-graph replay, no host sync; eager = "
                        "the restated reference update (6 .item() + float() read-backs, numpy get_state round trip)"}
    rows += [update_case()]

    # the Pvtol update (P sac_cbf_clf.py:181-331: 3 chained NODE predictions + 2 policy re-samples + get_obs + the 10-row
    # constraint kernel inside the actor loss): the same body launched eagerly through the fused kernels vs one graph replay
    def update_case_P():
        import numpy as np
        g = torch.load(os.path.join(ROOT, "tests", "golden", "update_P.pt"), weights_only=False)
        k, B = g["consts"], g["B"]
        C = nlbac_b200.constraints
        sp = type("S", (), {"low": np.array(k["action_low"], dtype="float32"), "high": np.array(k["action_high"], dtype="float32")})
        nets = {"critic": nlbac_b200.QNetwork(11, 2, 256), "critic_target": nlbac_b200.QNetwork(11, 2, 256),
                "lya": nlbac_b200.LyaNetwork(11, 256), "lya_t": nlbac_b200.LyaNetwork(11, 256),
                "policy": nlbac_b200.GaussianPolicy(11, 2, 256, sp), "backup": nlbac_b200.GaussianPolicy(11, 2, 256, sp),
                "node": nlbac_b200.NeuralODEModel(6, 6, 12)}
        for n, key in (("critic", "critic"), ("critic_target", "critic"), ("lya", "lyapunovNet"), ("lya_t", "lyapunovNet"),
                       ("policy", "policy"), ("backup", "backup_policy"), ("node", "neural_ode_model")):
            nets[n].load_state_dict(g["init"][key])
            nets[n].to(dev)
        kw = dict(dt=k["dt"], gamma_b=k["gamma_b"], hazard_locations=k["hazards"], hazards_radius=k["hazards_radius"],
                  safety_operator_follow=k["safety_operator_follow"], operator_dist=k["operator_dist"], y_max=k["y_max"],
                  y_min=k["y_min"], batch_size=B)
        node, lya = nets["node"], nets["lya"]
        dm = nlbac_b200.DynamicsModel(type("E", (), {"dynamics_mode": "Pvtol"}))
        def make(concurrent):
            h = {}
            u_ = nlbac_b200.GraphedUpdate(
                nets["critic"], nets["critic_target"], lya, nets["lya_t"], nets["policy"], nets["backup"], node, dm, B,
                primary=lambda s_, pi, V: C.pvtol_required_matrix(node, lya, nets["policy"], dm, s_, pi, V,
                                                                  eps_next=h["u"].eps[3], eps_next_next=h["u"].eps[4], **kw),
                backup=lambda s_, pi: C.pvtol_required_matrix(node, None, nets["backup"], dm, s_, pi, None,
                                                              eps_next=h["u"].eps[5], eps_next_next=h["u"].eps[6], **kw),
                num_constraints=10, obs_dim=11, act_dim=2, center_pos_num=11, ratio_floor=0.002, node_n_s=6, node_n_u=2,
                n_extra_eps=4, lya_on_obs=True, backup_update_interval=4, separate_backup_rho=True, concurrent=concurrent)
            h["u"] = u_
            return u_
        batch = tuple(None if x is None else x.to(dev) for x in g["batches"][0])
        upd1 = make(False)
        upd1.step(batch)
        o1_full = lat_gpu(lambda: upd1.step(batch, with_backup=True))
        o1_nob = lat_gpu(lambda: upd1.step(batch, with_backup=False))
        upd = make(True)
        upd.step(batch)
        o_full = lat_gpu(lambda: upd.step(batch, with_backup=True))
        o_nob = lat_gpu(lambda: upd.step(batch, with_backup=False))
        e_full = lat_gpu(lambda: upd._body(True), iters=20, warm=3)
        e_nob = lat_gpu(lambda: upd._body(False), iters=20, warm=3)
        return {"case": "config3-shaped Pvtol full update_parameters (3-step rollout in the actor loss): graph replay vs the "
                        "same fused launches issued eagerly", "batch": B, "ours_us": o_nob, "ours_with_backup_update_us": o_full,
                "ours_single_stream_graph_us": o1_nob, "ours_single_stream_graph_with_backup_us": o1_full,
                "ours_eager_launches_us": e_nob, "ours_eager_launches_with_backup_us": e_full, "torch_eager_gpu_us": None,
                "torch_cpu_us": None, "speedup_vs_eager_gpu": None, "speedup_vs_cpu": None,
                "graph_speedup_vs_own_eager": e_nob / o_nob}
    rows += [update_case_P()]

    # BASELINE config 3 as named: Pvtol with the learned barrier certificate (BP sac_cbf_clf.py), batch 256, the NODE
    # predictions inside the actor loss solved with adaptive dopri5 (device-side step controller -> capturable) vs euler
    def update_case_BP(solver):
        import numpy as np
        g = torch.load(os.path.join(ROOT, "tests", "golden", "update_BP.pt"), weights_only=False)
        k, B = g["consts"], g["B"]
        C = nlbac_b200.constraints
        sp = type("S", (), {"low": np.array(k["action_low"], dtype="float32"), "high": np.array(k["action_high"], dtype="float32")})
        nets = {"critic": nlbac_b200.QNetwork(11, 2, 256), "critic_target": nlbac_b200.QNetwork(11, 2, 256),
                "lya": nlbac_b200.LyaNetwork(11, 256), "lya_t": nlbac_b200.LyaNetwork(11, 256),
                "policy": nlbac_b200.GaussianPolicy(11, 2, 256, sp), "node": nlbac_b200.NeuralODEModel(6, 6, 12),
                "bar": nlbac_b200.BarrierNetwork(11, 2, 256), "bar_t": nlbac_b200.BarrierNetwork(11, 2, 256)}
        for n, key in (("critic", "critic"), ("critic_target", "critic"), ("lya", "lyapunovNet"), ("lya_t", "lyapunovNet"),
                       ("policy", "policy"), ("node", "neural_ode_model"), ("bar", "BarrierNet"), ("bar_t", "BarrierNet")):
            nets[n].load_state_dict(g["init"][key])
            nets[n].to(dev)
        node, lya = nets["node"], nets["lya"]
        dm = nlbac_b200.DynamicsModel(type("E", (), {"dynamics_mode": "Pvtol"}))
        kw = dict(dt=k["dt"], gamma_b=k["gamma_b"], safety_operator_follow=k["safety_operator_follow"], batch_size=B,
                  solver=solver)
        h = {}
        upd = nlbac_b200.GraphedUpdate(
            nets["critic"], nets["critic_target"], lya, nets["lya_t"], nets["policy"], None, node, dm, B,
            primary=lambda s_, pi, V: C.barrier_pvtol_required_matrix(
                node, lya, nets["bar"], nets["policy"], dm, h["u"].buf["state"], s_, pi, V, eps_next=h["u"].eps[3], **kw),
            backup=None, num_constraints=2, ratio_floor=0.002, n_extra_eps=1, barrier=nets["bar"],
            barrier_target=nets["bar_t"], lya_on_obs=True, node_n_s=6, node_n_u=2, obs_dim=11, act_dim=2,
            center_pos_num=11)
        h["u"] = upd
        batch = tuple(None if x is None else x.to(dev) for x in g["batches"][0])
        upd.step(batch)
        o = lat_gpu(lambda: upd.step(batch))
        e = lat_gpu(lambda: upd._body(False), iters=20, warm=3)
        return {"case": "config3 Pvtol + learned barrier full update_parameters, NODE solver = %s" % solver, "batch": B,
                "ours_us": o, "ours_eager_launches_us": e, "torch_eager_gpu_us": None, "torch_cpu_us": None,
                "speedup_vs_eager_gpu": None, "speedup_vs_cpu": None, "graph_speedup_vs_own_eager": e / o}
    rows += [update_case_BP("euler"), update_case_BP("dopri5")]

    # replay sampling: the reference's list-of-tuples memory + np.stack + H2D vs the device ring
    from oracle.replay import ReplayMemory as RefReplay
    import numpy as np
    rng = np.random.default_rng(0)
    trs = [(rng.normal(size=7), rng.normal(size=2), 0.1, 0.0, rng.normal(size=2), rng.normal(size=2),
            rng.normal(size=7), 1.0) for _ in range(40000)]
    ours_mem, ref_mem = nlbac_b200.ReplayMemory(40000, 0), RefReplay(40000, 0)
    for tr in trs:
        ours_mem.push(*tr)
        ref_mem.push(*tr)

    def ref_sample(bs):
        def f():
            batch = ref_mem.sample(bs)
            return [torch.FloatTensor(a).to(dev) for a in batch[:8]]
        return f
    for bs in (128, 32768):
        o = lat_gpu(lambda: ours_mem.sample(bs), iters=20, warm=3)
        c = lat_cpu(lambda: (ref_sample(bs)(), torch.cuda.synchronize()), iters=5, warm=1)
        rows.append({"case": "replay sample -> device batch", "batch": bs, "ours_us": o, "torch_eager_gpu_us": c,
                     "torch_cpu_us": c, "speedup_vs_eager_gpu": c / o, "speedup_vs_cpu": c / o})
    # vectorised environment stepping: E Unicycle environments per launch vs the numpy restatement of
    # the reference's env (one python object per environment in the reference itself: E x ~25 us)
    from oracle import envs as oenv
    for E in (1, 4096, 1 << 20):
        env = nlbac_b200.VecUnicycleEnv(E)
        act = torch.rand(E, 2, dtype=torch.float64, device=dev) * 2 - 1
        o = lat_gpu(lambda: env.step(act), iters=30, warm=5)
        st, last, stp = oenv.reset(E)
        a_np = act.cpu().numpy()
        c = lat_cpu(lambda: oenv.step(st, last, stp, a_np), iters=5, warm=1)
        rows.append({"case": "UnicycleEnv.step x E (float64)", "batch": E, "ours_us": o, "torch_eager_gpu_us": None,
                     "torch_cpu_us": c, "speedup_vs_eager_gpu": None, "speedup_vs_cpu": c / o,
                     "env_steps_per_s": E / (o * 1e-6), "hbm_gbs": E * 206 / (o * 1e-6) / 1e9})
    E = 1 << 20
    cenv, penv = nlbac_b200.VecSimulatedCarsEnv(E), nlbac_b200.VecPvtolEnv(E)
    ca = torch.rand(E, dtype=torch.float64, device=dev) * 2 - 1
    pa = torch.rand(E, 2, dtype=torch.float64, device=dev) * 2 - 1
    cst, ct, cstep = oenv.cars_reset(np.zeros(E))
    pst, pstep = oenv.pvtol_reset(E)
    ca_np, pa_np = ca.cpu().numpy(), pa.cpu().numpy()
    for name, fn, ref_fn, nbytes in (
            ("SimulatedCarsEnv.step x E (float64)", lambda: cenv.step(ca), lambda: oenv.cars_step(cst, ct, cstep, ca_np), 366),
            ("PvtolEnv.step x E (float64)", lambda: penv.step(pa), lambda: oenv.pvtol_step(pst, pstep, pa_np), 378)):
        o = lat_gpu(fn, iters=30, warm=5)
        c = lat_cpu(ref_fn, iters=3, warm=1)
        rows.append({"case": name, "batch": E, "ours_us": o, "torch_eager_gpu_us": None, "torch_cpu_us": c,
                     "speedup_vs_eager_gpu": None, "speedup_vs_cpu": c / o, "env_steps_per_s": E / (o * 1e-6),
                     "hbm_gbs": E * nbytes / (o * 1e-6) / 1e9})
    print(json.dumps({"suite": "configs", "unit": "microseconds per forward+backward call (median)", "cores": os.cpu_count(),
                      "note": "B = 128/256 calls are launch-latency bound: one fused launch per direction vs ~90 eager "
                              "launches (SURVEY.md §8 a3)", "rows": rows}))


if __name__ == "__main__":
    a = parse()
    if a.suite == "configs":
        run_configs(a)
    elif a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
