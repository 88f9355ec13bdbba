import sys, os; sys.path.insert(0,'tests'); sys.path.insert(0,'.')
import numpy as np, torch, bench
dev = torch.device('cuda:0')
w = dict(bench.WORKLOADS["c2"]); w["nb"] = 1_000_000
idx, queries, gt, cent_h = bench.build_index(w, torch, dev, lambda m: None)
Q = queries[:2000]
C = torch.from_numpy(cent_h).to(dev)
dd = (Q.double()[:, None, :] - C.double()[None, :, :])  if False else None
d2 = (Q.double() ** 2).sum(1)[:, None] - 2 * Q.double() @ C.double().T + (C.double() ** 2).sum(1)[None]
top = d2.topk(64, dim=1, largest=False)
res = {}
for impl in (0, 1):
    idx.set_option("coarse_impl", impl)
    cd, ci = idx.coarse(Q.cpu().numpy(), 64)
    res[impl] = (cd, ci)
    ti = top.indices.cpu().numpy()
    agree = np.mean([len(set(a) & set(b)) / 64 for a, b in zip(ci.tolist(), ti.tolist())])
    print("impl", impl, "set agreement with fp64 brute force", agree, "dist err", np.abs(cd - top.values.cpu().numpy()).max(), "sorted", np.all(np.diff(cd, axis=1) >= 0))
print("impl0 == impl1", np.mean(res[0][1] == res[1][1]))
r = np.nonzero((res[0][1] != res[1][1]).any(1))[0][:3]
for i in r:
    print(i, res[0][1][i][:8], res[1][1][i][:8], top.indices[i][:8].tolist())
    print("   ", res[0][0][i][:8], res[1][0][i][:8])
