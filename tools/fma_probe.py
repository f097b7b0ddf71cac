import sys, os, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, nlbac_b200
L = nlbac_b200._lib.lib()
sink = torch.zeros(4, device="cuda"); fl = ctypes.c_double(); st = torch.cuda.current_stream().cuda_stream
def t(fn):
    best = 0
    for i in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        if i: best = max(best, fl.value / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    return best
print("imm-fed FFMA peak  TF:", t(lambda: L.nlbac_fp32_peak_probe(4096, sink.data_ptr(), ctypes.byref(fl), st)))
print("reg-fed FFMA  8x12 TF:", t(lambda: L.nlbac_fma_shape_probe(0, 20000, sink.data_ptr(), ctypes.byref(fl), st)))
print("reg-fed FFMA2 8x12 TF:", t(lambda: L.nlbac_fma_shape_probe(1, 20000, sink.data_ptr(), ctypes.byref(fl), st)))
