import sys; sys.path.insert(0,'tests'); sys.path.insert(0,'.')
import numpy as np, torch, bench
dev = torch.device('cuda:0')
w = dict(bench.WORKLOADS["c2"]) if hasattr(bench, "WORKLOADS") else None
print({k: w[k] for k in ("nc", "d", "nlearn", "lloyd", "nsubc")})
P = bench.pkg()
means = bench.blob_means(w, torch, dev)
learn = bench.draw(w, torch, dev, means, w["nlearn"], 1234 + 2)
cent = learn[torch.randperm(w["nlearn"], device=dev, generator=torch.Generator(device=dev).manual_seed(7))[:w["nc"]]].clone()
idx = P.IndexIVF_HNSW_Grouping(w["d"], w["nc"], w["M"], 8, w["nsubc"])
idx.build_quantizer(centroids=cent.cpu().numpy())
x = bench.draw(w, torch, dev, means, 1_000_000, 99991)
res = {}
for impl in (0, 1):
    idx.set_option("coarse_impl", impl)
    lab = torch.empty(x.shape[0], dtype=torch.int32, device=dev); dt = torch.empty(x.shape[0], device=dev)
    idx.coarse_dev(x, 1, dt, lab); idx.sync()
    res[impl] = (lab.clone(), dt.clone())
    print("impl", impl, "stats", idx.stats() if hasattr(idx, "stats") else "")
a, b = res[0][0], res[1][0]
print("assign agree:", (a == b).float().mean().item(), "dist maxdiff", (res[0][1] - res[1][1]).abs().max().item())
bad = (a != b).nonzero().flatten()[:5]
print("bad rows", bad.tolist(), a[bad].tolist(), b[bad].tolist(), res[0][1][bad].tolist(), res[1][1][bad].tolist())
# brute force on first 50k
cn = (cent * cent).sum(1)
bf = (cn[None] - 2 * x[:50000] @ cent.T).argmin(1).int()
print("impl0 vs bf", (a[:50000] == bf).float().mean().item(), "impl1 vs bf", (b[:50000] == bf).float().mean().item())
# neighbour lists of the centroids (exclude self), both impls
nn = {}
for impl in (0, 1):
    j = P.IndexIVF_HNSW_Grouping(w["d"], w["nc"], w["M"], 8, w["nsubc"])
    j.set_option("coarse_impl", impl)
    j.build_quantizer(centroids=cent.cpu().numpy())
    rng = np.random.default_rng(0)
    j.set_pq(rng.standard_normal((w["M"], 256, w["d"] // w["M"])).astype(np.float32) * 0.05, np.sort(rng.random(256).astype(np.float32)), None)
    j.add_points_dev(x[:100000].contiguous(), torch.arange(100000, dtype=torch.int32, device=dev), None)
    j.compute_centroid_norms()
    m = j.export_index_dict()
    nn[impl] = {k: m[k] for k in ("nn_idx", "icd", "alphas", "sub_sizes", "list_off")}
    j.close()
print(type(nn[0]), [k for k in nn[0]] if isinstance(nn[0], dict) else None)
if isinstance(nn[0], dict):
    for k in nn[0]:
        a, b = np.asarray(nn[0][k]), np.asarray(nn[1][k])
        print(k, a.shape, "equal frac", np.mean(a == b))
    a, b = nn[0]["nn_idx"], nn[1]["nn_idx"]
    bad = np.nonzero((a != b).any(1))[0][:3]
    for r in bad:
        print(r, a[r][:10], b[r][:10], nn[0]["icd"][r][:5], nn[1]["icd"][r][:5])
    cc = cent
    selfin0 = np.mean((a == np.arange(a.shape[0])[:, None]).any(1)); selfin1 = np.mean((b == np.arange(a.shape[0])[:, None]).any(1))
    print("self present", selfin0, selfin1)
