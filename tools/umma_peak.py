"""Measured tcgen05 tensor-pipe peaks on this GPU (nlbac_umma_peak_probe): burst (one ~10 ms launch, best of 5) and
sustained (back-to-back launches for ~3 s, as MEASURED_PEAKS.json does for cuBLAS), per probe mode.  Prints JSON."""
import ctypes
import json
import os
import sys
import time

R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
import torch  # noqa: E402
import nlbac_b200  # noqa: E402

MODES = {0: "tf32_ss_n256", 1: "tf32_ts_n256", 2: "tf32_ss_n112", 3: "tf32_ts_n112", 4: "bf16_ss_n256"}


def measure(sustain_s=3.0, modes=(0, 1, 2, 3, 4)):
    L = nlbac_b200._lib.lib()
    st = torch.cuda.current_stream().cuda_stream
    out = {}
    for mode in modes:
        flops = ctypes.c_double()
        iters = 4000

        def launch():
            nlbac_b200._lib.check(L.nlbac_umma_peak_probe(mode, iters, ctypes.byref(flops), st), "umma_peak_probe")

        launch()
        torch.cuda.synchronize()
        best = 0.0
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); launch(); e1.record()
            torch.cuda.synchronize()
            best = max(best, flops.value / (e0.elapsed_time(e1) * 1e-3) / 1e12)
        res = {"burst_tflops": best}
        if sustain_s > 0:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            n, t0 = 0, time.perf_counter()
            e0.record()
            while time.perf_counter() - t0 < sustain_s:
                for _ in range(20):
                    launch()
                n += 20
                torch.cuda.synchronize()
            e1.record()
            torch.cuda.synchronize()
            res["sustained_tflops"] = n * flops.value / (e0.elapsed_time(e1) * 1e-3) / 1e12
        out[MODES[mode]] = res
    return out


if __name__ == "__main__":
    print(json.dumps({"umma_peak": measure(float(sys.argv[1]) if len(sys.argv) > 1 else 3.0)}))
