"""GPU parity at SCALE: indexes built on the GPU by the library itself (bench.build_index) at sizes the golden fixtures
cannot carry -- long lists and sub-groups (the kept-run tiling of the scan), OPQ, nprobe 32..64 -- checked (a) against
the plain-C port on a sample of queries and (b) through size-independent properties of the search (sortedness,
idempotence, k-prefix, sharded == unsharded, pruning only removes candidates)."""
import importlib
import os
import sys

import numpy as np
import pytest

from conftest import ROOT, load_pkg, dist_close, labels_match_up_to_ties

pytestmark = pytest.mark.gpu

# SIFT1B-shape in the small: ~1000-code lists, ~15-code sub-groups, OPQ, Grouping + pruning (BASELINE configs[3] scaled)
W_LONG = dict(kind="sift", nb=2_000_000, d=128, nc=2048, M=16, nsubc=64, pruning=True, opq=True, nprobe=32,
              max_codes=20000, k=10, nq=256, nlearn=200_000, ntrain=30000, lloyd=1, efs=64)
# DEEP-shape (configs[1]) at 1/5 scale: short lists, sub-groups of 2-3 codes (the scan walks whole parts)
W_DEEP = dict(kind="deep", nb=2_000_000, d=96, nc=16384, M=16, nsubc=64, pruning=True, opq=False, nprobe=64,
              max_codes=30000, k=10, nq=256, nlearn=400_000, ntrain=30000, lloyd=1, efs=64)


# shapes that take the other K-G1b paths: d = 104 uses 96 dims with a 104-float row pitch (fstdistfunc drops the
# d % 16 tail), d = 512 does not fit the shared-memory ring and falls back to one thread per row
W_D104 = dict(kind="sift", nb=200_000, d=104, nc=512, M=8, nsubc=32, pruning=True, opq=False, nprobe=16,
              max_codes=4000, k=10, nq=128, nlearn=60_000, ntrain=20000, lloyd=1, efs=64)
W_D512 = dict(kind="deep", nb=100_000, d=512, nc=256, M=8, nsubc=16, pruning=True, opq=False, nprobe=8,
              max_codes=3000, k=10, nq=128, nlearn=40_000, ntrain=20000, lloyd=1, efs=64)
# more than 64 sub-groups per list (the scan keeps their bases in shared memory instead of two registers) and 32-byte
# codes (one 32-wide window per warp step)
W_NS128 = dict(kind="sift", nb=200_000, d=64, nc=256, M=16, nsubc=128, pruning=True, opq=False, nprobe=12,
               max_codes=6000, k=10, nq=128, nlearn=60_000, ntrain=20000, lloyd=1, efs=64)
W_M32 = dict(kind="deep", nb=300_000, d=96, nc=512, M=32, nsubc=64, pruning=True, opq=True, nprobe=24,
             max_codes=8000, k=10, nq=128, nlearn=60_000, ntrain=20000, lloyd=1, efs=64)
# IVFADC (configs[0]-like): long lists walked whole, max_codes cutting the probe sequence
W_ADC = dict(kind="sift", nb=1_000_000, d=128, nc=1024, M=8, nsubc=0, pruning=False, opq=False, nprobe=32,
             max_codes=12000, k=10, nq=128, nlearn=100_000, ntrain=20000, lloyd=1, efs=64)
WORKLOADS = {"adc": W_ADC, "long_opq": W_LONG, "deep": W_DEEP, "d104": W_D104, "d512": W_D512, "nsubc128": W_NS128, "m32": W_M32}


def _bench():
    sys.path.insert(0, ROOT)
    return importlib.import_module("bench")


@pytest.fixture(scope="module", params=list(WORKLOADS))
def built(request):
    import torch
    bench = _bench()
    w = dict(WORKLOADS[request.param])
    dev = torch.device("cuda:0")
    idx, queries, gt, cent_h = bench.build_index(w, torch, dev, lambda m: None)
    yield w, idx, queries, gt, cent_h
    idx.close()


def test_cuda_equals_port_on_gpu_built_index(port, built):
    """Same candidate lists in -> labels identical up to equal-distance ties, distances within 1e-4 relative, pruning
    thresholds to 1e-5, scanned-code counts EXACTLY (BASELINE.json north_star)."""
    w, idx, queries, gt, cent_h = built
    k, nq = w["k"], 64
    Q = queries[:nq].cpu().numpy()
    idx.set_option("debug", 1)
    cd, ci = idx.coarse(Q, w["nprobe"])
    D, L = idx.search2(k, Q, cd, ci)
    thr, scanned = idx.last_thresholds(nq)
    idx.set_option("debug", 0)
    ix = idx.export_index_dict()
    ix.update(centroids=idx.get_centroids(), pqc=idx.pq_centroids, norm_table=idx.norm_centroids,
              opq=idx.opq_matrix if w["opq"] else None)
    Dp, Lp, pthr, pscanned, pncode = port.search(ix, Q, cd, ci, k, w["nprobe"], w["max_codes"], w["pruning"], debug=True)
    o = np.argsort(Dp, axis=1, kind="stable")
    Dps, Lps = np.take_along_axis(Dp, o, 1), np.take_along_axis(Lp, o, 1)
    scale = float(ix["centroid_norms"].max())
    assert dist_close(D, Dps, scale)
    assert labels_match_up_to_ties(D, L, Dps, Lps, scale) == 0
    # position-wise equality is only expected where distances are distinct (8-byte codes of 512-d vectors collide)
    assert np.mean(L == Lps) >= (0.99 if w["M"] >= 16 else 0.9)
    assert np.array_equal(scanned.astype(np.int64), np.asarray(pncode).astype(np.int64))
    if w["nsubc"] and w["pruning"]:
        assert np.allclose(thr, pthr, rtol=1e-5, equal_nan=True)   # no probed code at all: 0/0 on both sides


def test_search_properties_at_scale(built):
    import torch
    w, idx, queries, gt, cent_h = built
    sharding = importlib.import_module("ivf-hnsw_b200.sharding")
    dev = queries.device
    nq, k = queries.shape[0], w["k"]
    D = torch.empty(nq, k, device=dev)
    L = torch.empty(nq, k, dtype=torch.int64, device=dev)
    idx.search_dev(k, queries, D, L)
    torch.cuda.synchronize()
    # sorted ascending, labels valid and distinct per query
    assert bool((D[:, 1:] >= D[:, :-1]).all())
    assert bool((L >= 0).all()) and bool((L < w["nb"]).all())
    assert all(len(set(r)) == k for r in L.cpu().tolist())
    # idempotent
    D2 = torch.empty_like(D)
    L2 = torch.empty_like(L)
    idx.search_dev(k, queries, D2, L2)
    torch.cuda.synchronize()
    assert torch.equal(D, D2) and torch.equal(L, L2)
    # the k best are a prefix of the 100 best (k <= 32 and k <= 128 take different result-list paths)
    D100 = torch.empty(nq, 100, device=dev)
    L100 = torch.empty(nq, 100, dtype=torch.int64, device=dev)
    idx.search_dev(100, queries, D100, L100)
    torch.cuda.synchronize()
    assert torch.equal(D, D100[:, :k]) and torch.equal(L, L100[:, :k])
    # ... and the 100 best of the 200 best (k > 128: the locked shared-memory result list)
    D200 = torch.empty(nq, 200, device=dev)
    L200 = torch.empty(nq, 200, dtype=torch.int64, device=dev)
    idx.search_dev(200, queries, D200, L200)
    torch.cuda.synchronize()
    assert torch.equal(D100, D200[:, :100]) and torch.equal(L100, L200[:, :100])
    # recall of the true nearest neighbour is far above chance
    if w["M"] >= 16:      # (8-byte codes of 100+-dimensional vectors are too coarse for a meaningful bound)
        assert float((L100 == gt[:nq, None]).any(1).float().mean()) > 0.3
    # sharded by centroid range (3 shards, query-sliced preparation) == unsharded, bit for bit
    R = 3
    per = (nq + R - 1) // R
    tables = torch.zeros(R * per, idx.tables_bytes(), dtype=torch.uint8, device=dev)
    for r in range(R):
        b, e, _ = sharding.query_slice(r, R, nq)
        idx.prepare_dev(queries, b, e, tables)
    gD = torch.empty(R, nq, k, device=dev)
    gS = torch.empty(R, nq, k, dtype=torch.int32, device=dev)
    gL = torch.empty(R, nq, k, dtype=torch.int64, device=dev)
    for r in range(R):
        idx.set_shard(*sharding.shard_range(r, R, w["nc"]))
        idx.scan_tables_dev(k, queries, tables, gD[r], gS[r], gL[r])
    idx.set_shard(0, w["nc"])
    D3 = torch.empty_like(D)
    L3 = torch.empty_like(L)
    idx.merge_partials_dev(k, nq, R, gD, gS, gL, D3, L3)
    torch.cuda.synchronize()
    assert torch.equal(D, D3) and torch.equal(L, L3)
    # pruning only removes candidates: without it (and without the max_codes cut) no result gets worse
    idx.do_pruning, idx.max_codes = False, 10 ** 9
    Dn = torch.empty_like(D)
    Ln = torch.empty_like(L)
    idx.search_dev(k, queries, Dn, Ln)
    torch.cuda.synchronize()
    idx.do_pruning, idx.max_codes = w["pruning"], w["max_codes"]
    assert bool((Dn <= D).all())
