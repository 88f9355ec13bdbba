import sys, os; sys.path.insert(0,'tests'); sys.path.insert(0,'.')
import numpy as np, torch, bench
dev = torch.device('cuda:0')
w = dict(bench.WORKLOADS["c2"]); w["nb"] = 1_000_000
idx, queries, gt, cent_h = bench.build_index(w, torch, dev, lambda m: None)
means = bench.blob_means(w, torch, dev)
C = torch.from_numpy(cent_h).to(dev); cn = (C * C).sum(1)
for b0 in range(0, 10_000_000, 1_000_000):
    x = bench.draw(w, torch, dev, means, 1_000_000, 99991 + b0)
    res = {}
    for impl in (0, 1):
        idx.set_option("coarse_impl", impl)
        lab = torch.empty(x.shape[0], dtype=torch.int32, device=dev); dt = torch.empty(x.shape[0], device=dev)
        idx.coarse_dev(x, 1, dt, lab); idx.sync()
        res[impl] = lab.long()
    bf = (cn[None] - 2 * x[:20000] @ C.T).argmin(1)
    print(b0, "impl0==impl1", (res[0] == res[1]).float().mean().item(), "impl0 vs bf", (res[0][:20000] == bf).float().mean().item(),
          "impl1 vs bf", (res[1][:20000] == bf).float().mean().item(), flush=True)
