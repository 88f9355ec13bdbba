import sys; sys.path.insert(0,'tests'); sys.path.insert(0,'.')
import numpy as np, torch
from conftest import load_pkg
pkg = load_pkg()
d, nc, n = 96, 65536, 200000
g = torch.Generator(device='cuda').manual_seed(1)
C = torch.randn(nc, d, device='cuda', generator=g) * 0.3
X = C[torch.randint(0, nc, (n,), device='cuda', generator=g)] + 0.2 * torch.randn(n, d, device='cuda', generator=g)
idx = pkg.IndexIVF_HNSW(d, nc, 16, 8)
idx.build_quantizer(centroids=C.cpu().numpy())
Xh = X.cpu().numpy()
best = torch.empty(n, dtype=torch.long, device='cuda')
cn = (C * C).sum(1)
for i in range(0, n, 20000):
    dd = cn[None, :] - 2 * X[i:i+20000] @ C.T
    best[i:i+20000] = dd.argmin(1)
best = best.cpu().numpy()
for impl in (0, 1):
    idx.set_option("coarse_impl", impl)
    lab = idx.assign(Xh)
    print("impl", impl, "agree with brute force:", np.mean(lab == best))
    cd, ci = idx.coarse(Xh[:2000], 64)
    dd = (cn[None, :] - 2 * X[:2000] @ C.T)
    top = dd.topk(64, dim=1, largest=False).indices.cpu().numpy()
    print("   nprobe=64 set agreement:", np.mean([len(set(a) & set(b)) / 64 for a, b in zip(ci.tolist(), top.tolist())]))
