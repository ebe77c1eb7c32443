import torch, sys
sys.path.insert(0, '.')
from b200net import ops
dev = torch.device('cuda:0')
g = torch.Generator().manual_seed(1)
rows, H, p = 512, 768, 0.1
x = (torch.randn((rows, H), generator=g).to(torch.bfloat16).to(dev)) + 3.0
zero = torch.zeros_like(x)
gamma = torch.ones(H, device=dev); beta = torch.zeros(H, device=dev)
s, y, mean, rstd = ops.ln_fwd(x.clone(), zero, gamma, beta, 1e-12, p=p, seed=1234, offset=77 << 34, inplace_sum=False)
kept = s != 0
g = torch.Generator().manual_seed(5)
dy = (torch.randn((rows, H), generator=g).to(torch.bfloat16).to(dev)) + 3.0
dres, dx, dg, db, dbias = ops.ln_bwd(dy, None, s, mean, rstd, gamma, p=p, seed=1234, offset=77 << 34)
a = dx != 0
b = kept & (dres != 0)
mm = a != b
print("mismatches", mm.sum().item(), "a&~b", (a & ~b).sum().item(), "~a&b", (~a & b).sum().item())
idx = mm.nonzero()[:10]
for r, c in idx.tolist():
    print(r, c, "x", x[r, c].item(), "s", s[r, c].item(), "dres", dres[r, c].item(), "dx", dx[r, c].item(), "kept", kept[r, c].item())
