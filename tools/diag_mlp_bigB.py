import sys; sys.path.insert(0, "/root/repo")
import torch, nlbac_b200
from oracle import nlbac_ref
def rel(a,b):
    a,b=a.detach().double().cpu(),b.detach().double().cpu(); return ((a-b).abs().max()/b.abs().max()).item()
for B in (4096, 70001, 300000):
    torch.manual_seed(B)
    ref = nlbac_ref.LyaNetwork(11, 256)
    with torch.no_grad():
        for p in ref.parameters():
            if p.ndim == 1: p.normal_(0, 0.1)
    ours = nlbac_b200.LyaNetwork(11, 256); ours.load_state_dict(ref.state_dict()); ours.cuda()
    x = torch.randn(B, 11); cot = torch.randn(B, 1)
    r64 = nlbac_ref.LyaNetwork(11,256).double(); r64.load_state_dict({k:v.double() for k,v in ref.state_dict().items()})
    (r64(x.double()) * cot.double()).sum().backward()
    (ref(x) * cot).sum().backward()
    xo = x.cuda(); (ours(xo) * cot.cuda()).sum().backward()
    for (k,p),(_,r),(_,d) in zip(ours.named_parameters(), ref.named_parameters(), r64.named_parameters()):
        print(B, k, "ours-vs-f32 %.1e  ours-vs-f64 %.1e  f32-vs-f64 %.1e" % (rel(p.grad, r.grad), rel(p.grad, d.grad), rel(r.grad, d.grad)))
