import ctypes, itertools, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, nlbac_b200
L = nlbac_b200._lib.lib()
L.nlbac_umma_probe.argtypes = [ctypes.POINTER(ctypes.c_int), ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
g = torch.Generator().manual_seed(0)
X = torch.randn(128, 104, generator=g); W = torch.randn(104, 104, generator=g); Y = torch.randn(128, 104, generator=g)
def run(p, Xc, Yc):
    D = torch.full((128, 112), float("nan"), device="cuda")
    arr = (ctypes.c_int * 10)(*p)
    L.nlbac_umma_probe(arr, Xc.data_ptr(), Yc.data_ptr(), D.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return D.cpu().double()
def err(got, want):
    g = got[:want.shape[0], :want.shape[1]]
    return ((g - want).abs().max() / want.abs().max()).item(), int((g != 0).sum()), bool(torch.isnan(g).any())
XCH, WCH = 2048, 1792
want1 = X.double() @ W.double()            # D[t][k] = sum_j X[t][j] W[j][k]
want2 = X.double().t() @ Y.double()        # D[j][k] = sum_t X[t][j] Y[t][k]
print("mode1: A K-major ok; B MN-major variants (lboB, sboB, stepB)")
for lbo, sbo, step in [(128, WCH, 128), (WCH, 128, 128), (128, WCH, 256), (WCH, 128, 256), (128, WCH, 64), (WCH,128,64)]:
    print((lbo, sbo, step), err(run([0, 1, XCH, 128, 2 * XCH, lbo, sbo, step, 13, 1], X.cuda(), W.cuda()), want1))
print("mode2: both MN-major variants")
for lbo, sbo, step in [(128, XCH, 128), (XCH, 128, 128), (128, XCH, 256), (XCH, 128, 256)]:
    print((lbo, sbo, step), err(run([1, 1, lbo, sbo, step, lbo, sbo, step, 16 if step == 128 else 8, 0], X.cuda(), Y.cuda()), want2))

print("mode3: bf16 both MN-major interleave (lbo, sbo, step) K=16 per MMA")
Xb, Yb = X.bfloat16().double(), Y.bfloat16().double()
want3 = Xb.t() @ Yb
for lbo, sbo, step in [(128, XCH, 256), (XCH, 128, 256)]:
    print((lbo, sbo, step), err(run([1, 1, lbo, sbo, step, lbo, sbo, step, 8, 2], X.cuda(), Y.cuda()), want3))
