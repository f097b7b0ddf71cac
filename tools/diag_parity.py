"""Diagnostic: per-tensor errors of the CUDA path vs fp64 oracle, next to the fp32 CPU oracle's own error."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
import test_node_gpu as T

def rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    d = (a - b).abs()
    return (d.max() / b.abs().max()).item(), d

for tag, B, method, t in [("Q", 9500, "rk4", torch.tensor([0.0, 0.02, 0.04])), ("U", 256 * 148 + 37, "rk4", torch.tensor([0.0, 0.02])),
                          ("U", 4000, "rk4", torch.tensor([0.0, 0.02])), ("U", 128, "euler", torch.tensor([0.0, 0.02]))]:
    got, outs = T._run_pair(tag, B, method, t)
    names = ["sol", "gy0"] + ["p%d" % i for i in range(len(got[2]))]
    G = [got[0], got[1]] + list(got[2]); O32 = [outs["f32"][0], outs["f32"][1]] + list(outs["f32"][2]); O64 = [outs["f64"][0], outs["f64"][1]] + list(outs["f64"][2])
    print("==", tag, B, method)
    for n, a, b32, b64 in zip(names, G, O32, O64):
        e, d = rel(a, b64); o, d2 = rel(b32, b64)
        # concentration: how many entries exceed 10% of the max error
        cnt = int((d > 0.1 * d.max()).sum()); tot = d.numel()
        where = ""
        if n in ("sol", "gy0"):
            idx = d.reshape(-1, d.shape[-1]).max(1).values.argmax().item(); where = "row %d" % idx
        print("  %-4s ours %.2e  cpu32 %.2e  (entries >10%% of max err: %d/%d) %s" % (n, e, o, cnt, tot, where))
