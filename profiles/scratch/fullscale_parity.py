import sys, time; sys.path.insert(0,'tests'); sys.path.insert(0,'.'); sys.path.insert(0,'oracle')
import numpy as np, torch, bench
import oracle as O
dev = torch.device('cuda:0')
w = dict(bench.WORKLOADS["c2"]); w["nb"] = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
idx, queries, gt, cent_h = bench.build_index(w, torch, dev, lambda m: print("[build]", m, flush=True))
k, nq = 10, 64
Q = queries[:nq].cpu().numpy()
idx.set_option("debug", 1)
D, L = idx.search(k, Q)
thr, scanned = idx.last_thresholds(nq)
cd, ci = idx.coarse(Q, w["nprobe"])
print("scanned/query (gpu)", scanned.mean(), "thr mean", thr.mean())
ix = idx.export_index_dict()
ix.update(centroids=cent_h, pqc=idx.pq_centroids, norm_table=idx.norm_centroids, opq=None)
port = O.Port()
t0 = time.time()
Dp, Lp, pthr, pscanned, pncode = port.search(ix, Q, cd, ci, k, w["nprobe"], w["max_codes"], w["pruning"], debug=True)
print("port %.1fs" % (time.time() - t0), "scanned/query (port)", np.mean(pncode), "thr mean", pthr.mean())
o = np.argsort(Dp, axis=1, kind="stable"); Dps, Lps = np.take_along_axis(Dp, o, 1), np.take_along_axis(Lp, o, 1)
print("labels equal frac", np.mean(L == Lps), "max rel dist err", np.max(np.abs(D - Dps) / np.maximum(np.abs(Dps), 1e-6)))
print("thr allclose", np.allclose(thr, pthr, rtol=1e-5), "scanned equal", np.array_equal(scanned.astype(np.int64), np.asarray(pncode).astype(np.int64)))
