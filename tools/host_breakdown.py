"""Wall-clock breakdown of a launch-bound NODE call (B = 128 Unicycle euler) on the host side."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, nlbac_b200
m = nlbac_b200.NeuralODEModel(3, 3, 6).cuda()
y = torch.rand(128, 5, device="cuda"); ps = list(m.parameters()); t = [0.0, 0.02]
def timeit(fn, n=300):
    for _ in range(30): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e6
def f_nograd():
    with torch.no_grad(): nlbac_b200.odeint(m, y, t, method="euler")
def f_fwd(): nlbac_b200.odeint(m, y, t, method="euler")
def f_fwd_sum(): nlbac_b200.odeint(m, y, t, method="euler")[-1][:, :3].sum()
def f_all():
    torch.autograd.grad(nlbac_b200.odeint(m, y, t, method="euler")[-1][:, :3].sum(), ps)
def f_all_bw():
    for p in ps: p.grad = None
    nlbac_b200.odeint(m, y, t, method="euler")[-1][:, :3].sum().backward()
lin = torch.nn.Linear(5, 3).cuda()
def f_lin():
    torch.autograd.grad(lin(y).sum(), list(lin.parameters()))
print("no_grad fwd      %.1f us" % timeit(f_nograd))
print("fwd (graph)      %.1f us" % timeit(f_fwd))
print("fwd + slice/sum  %.1f us" % timeit(f_fwd_sum))
print("fwd + autograd.grad %.1f us" % timeit(f_all))
print("fwd + .backward()   %.1f us" % timeit(f_all_bw))
print("torch Linear fwd+grad (floor) %.1f us" % timeit(f_lin))
import ctypes
L = nlbac_b200._lib.lib()
L.nlbac_kernel_timing(1)
for _ in range(100): f_all()
ms, n = (ctypes.c_double * 4)(), (ctypes.c_int * 4)()
L.nlbac_kernel_timing_read(ms, n); L.nlbac_kernel_timing(0)
print("GPU us per call: fwd %.1f  bwd sweep %.1f  fold %.1f   (launches %s)" % (ms[0] * 10, ms[1] * 10, ms[3] * 10, list(n)))
for B in (256, 1024, 32768):
    y = torch.rand(B, 5, device="cuda")
    L.nlbac_kernel_timing(1)
    for _ in range(50): f_all()
    L.nlbac_kernel_timing_read(ms, n); L.nlbac_kernel_timing(0)
    print("B=%d GPU us per call: fwd %.1f  bwd sweep %.1f  fold %.1f" % (B, ms[0] * 20, ms[1] * 20, ms[3] * 20))
