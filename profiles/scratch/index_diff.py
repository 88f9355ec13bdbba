import sys, os; sys.path.insert(0,'tests'); sys.path.insert(0,'.')
import numpy as np, torch, bench
dev = torch.device('cuda:0')
w = dict(bench.WORKLOADS["c2"]); w["nb"] = int(sys.argv[1]) if len(sys.argv) > 1 else 3_000_000
out = {}
for impl in (0, 1):
    os.environ["IVFHNSW_B200_COARSE_IMPL"] = str(impl)
    idx, queries, gt, cent_h = bench.build_index(w, torch, dev, lambda m: None)
    ix = idx.export_index_dict()
    ix["cent"] = cent_h; ix["pqc"] = idx.pq_centroids
    out[impl] = ix
    idx.close()
a, b = out[0], out[1]
for k in ("cent", "pqc", "list_off", "ids", "nn_idx", "icd", "alphas", "sub_sizes", "codes", "norm_codes", "centroid_norms"):
    x, y = np.asarray(a[k]), np.asarray(b[k])
    if x.shape != y.shape: print(k, "SHAPE", x.shape, y.shape); continue
    eq = np.mean(x == y)
    extra = "" if x.dtype.kind != "f" else " maxabs %.3g" % np.max(np.abs(x.astype(np.float64) - y.astype(np.float64)))
    print(k, x.shape, "equal frac %.6f" % eq, extra)
sa, sb = np.diff(a["list_off"]), np.diff(b["list_off"])
print("list size mean/max", sa.mean(), sa.max(), sb.mean(), sb.max(), "n lists differing", np.sum(sa != sb))
