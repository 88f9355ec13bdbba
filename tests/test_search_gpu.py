"""GPU parity tests: the CUDA path (through the C ABI) against the golden fixtures of the compiled
reference and against the plain-C port, on the reference's own candidate lists (search2 seam)."""
import importlib

import numpy as np
import pytest

from conftest import golden_names, load_golden, load_pkg, dist_close, labels_match_up_to_ties

pytestmark = pytest.mark.gpu


def make_index(pkg, ix, search_time=True):
    cls = pkg.IndexIVF_HNSW_Grouping if ix["nsubc"] else pkg.IndexIVF_HNSW
    args = (ix["d"], ix["nc"], ix["M"], 8) + ((ix["nsubc"],) if ix["nsubc"] else ())
    idx = cls(*args)
    idx.build_quantizer(centroids=ix["centroids"] if search_time else ix["centroids_build"])
    idx.set_pq(ix["pqc"], ix["norm_table"], ix["opq"])
    idx.do_opq = ix["opq"] is not None
    return idx


@pytest.mark.parametrize("name", golden_names())
def test_search_given_reference_candidates(name):
    """Identical coarse candidate lists in -> labels identical up to equal-distance ties, distances
    within 1e-4 relative (BASELINE.json north_star)."""
    pkg = load_pkg()
    g, ix, runs = load_golden(name)
    k, nprobe = int(g["k"]), int(g["nprobe"])
    scale = float(ix["centroid_norms"].max())
    idx = make_index(pkg, ix)
    idx.load_index_dict(ix)
    idx.nprobe = nprobe
    for tag, mc, pr in runs:
        idx.max_codes, idx.do_pruning = mc, pr
        D, L = idx.search2(k, g["Q"], g["cd_" + tag], g["ci_" + tag])
        Dr, Lr = g["D_" + tag], g["L_" + tag]
        o = np.argsort(Dr, axis=1, kind="stable")
        Drs, Lrs = np.take_along_axis(Dr, o, 1), np.take_along_axis(Lr, o, 1)
        assert dist_close(D, Drs, scale), (name, tag, np.abs(D - Drs).max())
        assert labels_match_up_to_ties(D, L, Drs, Lrs, scale) == 0, (name, tag)
        exact = np.mean([set(a) == set(b) for a, b in zip(L.tolist(), Lr.tolist())])
        assert exact >= 0.95, (name, tag, exact)
    idx.close()


@pytest.mark.parametrize("name", golden_names())
@pytest.mark.parametrize("split", [1, 3])
def test_search_splits_and_port(port, name, split):
    """Same answers whatever the split factor (partial top-k + merge), checked against the port;
    Grouping additionally checks the pruning threshold and the number of scanned codes."""
    pkg = load_pkg()
    g, ix, runs = load_golden(name)
    k, nprobe = int(g["k"]), int(g["nprobe"])
    scale = float(ix["centroid_norms"].max())
    idx = make_index(pkg, ix)
    idx.load_index_dict(ix)
    idx.nprobe = nprobe
    idx.set_option("scan_split", split)
    idx.set_option("debug", 1)
    for tag, mc, pr in runs:
        idx.max_codes, idx.do_pruning = mc, pr
        D, L = idx.search2(k, g["Q"], g["cd_" + tag], g["ci_" + tag])
        Dp, Lp, thr, scanned, ncode = port.search(ix, g["Q"], g["cd_" + tag], g["ci_" + tag], k, nprobe, mc, pr,
                                                  debug=True)
        o = np.argsort(Dp, axis=1, kind="stable")
        Dps, Lps = np.take_along_axis(Dp, o, 1), np.take_along_axis(Lp, o, 1)
        assert dist_close(D, Dps, scale), (name, tag)
        assert labels_match_up_to_ties(D, L, Dps, Lps, scale) == 0, (name, tag)
        gthr, gsc = idx.last_thresholds(D.shape[0])
        assert np.array_equal(gsc.astype(np.uint64), ncode), (name, tag)
        if ix["nsubc"] and pr:
            assert np.allclose(gthr, thr, rtol=1e-5), (name, tag)
    idx.close()


@pytest.mark.parametrize("name", golden_names())
@pytest.mark.parametrize("impl", [0, 1, 2, 3])
def test_coarse_exact_matches_port(port, name, impl):
    """K-A2: exact nprobe nearest centroids, ascending (dist, idx), fstdistfunc numerics; impl 0 is the
    fp32 SIMT GEMM, impl 1 the tcgen05 split-fp16 GEMM with three products, impl 2 the same with two products and
    a widened select threshold, impl 3 (default) ONE product with integer keys -- all feed the same exact re-rank."""
    pkg = load_pkg()
    g, ix, runs = load_golden(name)
    nprobe = int(g["nprobe"])
    idx = make_index(pkg, ix)
    idx.set_option("coarse_impl", impl)
    cd, ci = idx.coarse(g["Q"], nprobe)
    Q = g["Q"] if ix["opq"] is None else port.opq_apply(ix["opq"], g["Q"])
    pd, pi = port.coarse(Q, ix["centroids"], nprobe)
    assert np.allclose(cd, pd, rtol=1e-5, atol=1e-4)
    agree = np.mean(ci == pi)
    assert agree >= 0.999, agree       # only fp-tie swaps may differ
    assert np.all(np.diff(cd, axis=1) >= 0)
    idx.close()


@pytest.mark.parametrize("name", golden_names())
def test_full_search_recall_not_below_reference(port, name):
    """Full GPU path (exact coarse + scan): every returned distance is a true ADC distance of the
    port on the GPU's own candidates, and 1-recall@k vs the reference's answers is >= what the
    reference's HNSW coarse stage achieved against the exhaustive-probe answer."""
    pkg = load_pkg()
    g, ix, runs = load_golden(name)
    k, nprobe = int(g["k"]), int(g["nprobe"])
    scale = float(ix["centroid_norms"].max())
    idx = make_index(pkg, ix)
    idx.load_index_dict(ix)
    idx.nprobe = nprobe
    tag, mc, pr = runs[0]
    idx.max_codes, idx.do_pruning = mc, pr
    D, L = idx.search(k, g["Q"])
    cd, ci = idx.coarse(g["Q"], nprobe)
    Dp, Lp = port.search(ix, g["Q"], cd, ci, k, nprobe, mc, pr)
    o = np.argsort(Dp, axis=1, kind="stable")
    Dps, Lps = np.take_along_axis(Dp, o, 1), np.take_along_axis(Lp, o, 1)
    assert dist_close(D, Dps, scale)
    assert labels_match_up_to_ties(D, L, Dps, Lps, scale) == 0
    # best distance found by the exact coarse stage is never worse than the HNSW one
    Dr = np.sort(g["D_" + tag], axis=1)
    assert np.mean(D[:, 0] <= Dr[:, 0] * (1 + 1e-4) + 1e-3 * scale * 1e-2) >= 0.99
    idx.close()


@pytest.mark.parametrize("shape", [(96, 1000, 300, 16), (128, 4097, 513, 32), (32, 130, 7, 8), (200, 2500, 260, 20),
                                   (104, 777, 129, 5)])
@pytest.mark.parametrize("spread", [1.0, 300.0])
def test_coarse_tensor_core_vs_brute_force(shape, spread):
    """tcgen05 coarse GEMM on ragged shapes (d not a multiple of 64, nc not a multiple of 128, nq tails,
    large dynamic range): same neighbours as a float64 brute force."""
    pkg = load_pkg()
    d, nc, nq, nprobe = shape
    rng = np.random.default_rng(d * 7 + nc)
    C = (rng.standard_normal((nc, d)) * spread).astype(np.float32)
    Q = (C[rng.integers(0, nc, nq)] + 0.3 * spread * rng.standard_normal((nq, d))).astype(np.float32)
    Q[0] = 0
    idx = pkg.IndexIVF_HNSW(d, nc, 8, 8)
    idx.build_quantizer(centroids=C)
    res = {}
    for impl in (0, 1, 2, 3):
        idx.set_option("coarse_impl", impl)
        res[impl] = idx.coarse(Q, nprobe)
    dd = ((Q.astype(np.float64)[:, None, :(d // 16) * 16] - C.astype(np.float64)[None, :, :(d // 16) * 16]) ** 2).sum(-1)
    want = np.sort(dd, axis=1)[:, :nprobe]
    for impl in (0, 1, 2, 3):
        cd, ci = res[impl]
        assert np.allclose(cd, want, rtol=2e-5, atol=1e-5 * spread * spread), impl
        got = np.take_along_axis(dd, ci.astype(np.int64), 1)
        assert np.allclose(got, want, rtol=2e-5, atol=1e-5 * spread * spread), impl
    assert np.mean(res[0][1] == res[1][1]) >= 0.999
    assert np.mean(res[0][1] == res[2][1]) >= 0.999
    assert np.mean(res[0][1] == res[3][1]) >= 0.999
    idx.close()


def test_empty_and_ragged_inputs(port):
    """Edge cases: nq = 0, k larger than the number of scanned codes (-1 / FLT_MAX padding like
    faiss::maxheap_heapify), empty probed lists, nprobe = 1."""
    pkg = load_pkg()
    g, ix, runs = load_golden("adc_d32")
    idx = make_index(pkg, ix)
    idx.load_index_dict(ix)
    idx.nprobe, idx.max_codes = 1, 10 ** 9
    D, L = idx.search(10, np.zeros((0, ix["d"]), np.float32))
    assert D.shape == (0, 10)
    k = 400   # more than one list holds
    cd, ci = g["cd_mc0_pr0"][:, :1].copy(), g["ci_mc0_pr0"][:, :1].copy()
    D, L = idx.search2(k, g["Q"], cd, ci)
    Dp, Lp = port.search(ix, g["Q"], cd, ci, k, 1, 10 ** 9)
    sizes = np.diff(ix["list_off"])[ci[:, 0]]
    for i in range(D.shape[0]):
        n = int(min(sizes[i], k))
        assert np.all(L[i, n:] == -1) and np.all(D[i, n:] > 1e38)
        assert set(L[i, :n].tolist()) == set(Lp[i][Lp[i] >= 0].tolist())
    idx.close()


@pytest.mark.parametrize("name", ["adc_sift_m16", "grp_sift_m16", "grp_deep_opq", "grp_long_m16"])
@pytest.mark.parametrize("nshards", [2, 3])
def test_sharded_search_equals_single_gpu(name, nshards):
    """SURVEY.md 8(e) on one GPU: query-sliced preparation into shard-independent records, per-shard scans of
    the owned centroid ranges, K-M1 merge by (distance, scan order) == the unsharded answer, bit for bit."""
    import importlib
    import torch
    pkg = load_pkg()
    sharding = importlib.import_module("ivf-hnsw_b200.sharding")
    g, ix, runs = load_golden(name)
    k, nprobe = int(g["k"]), int(g["nprobe"])
    idx = make_index(pkg, ix)
    idx.load_index_dict(ix)
    idx.nprobe = nprobe
    tag, mc, pr = runs[-1]
    idx.max_codes, idx.do_pruning = mc, pr
    dev = torch.device("cuda:0")
    Q = torch.from_numpy(np.ascontiguousarray(g["Q"])).to(dev)
    nq = Q.shape[0]
    D0 = torch.empty(nq, k, device=dev); L0 = torch.empty(nq, k, dtype=torch.int64, device=dev)
    idx.search_dev(k, Q, D0, L0)
    # records prepared slice by slice, as the ranks would
    per = (nq + nshards - 1) // nshards
    tables = torch.zeros(nshards * per, idx.tables_bytes(), dtype=torch.uint8, device=dev)
    for r in range(nshards):
        b, e, _ = sharding.query_slice(r, nshards, nq)
        idx.prepare_dev(Q, b, e, tables)
    gD = torch.empty(nshards, nq, k, device=dev); gS = torch.empty(nshards, nq, k, dtype=torch.int32, device=dev)
    gL = torch.empty(nshards, nq, k, dtype=torch.int64, device=dev)
    for r in range(nshards):
        idx.set_shard(*sharding.shard_range(r, nshards, ix["nc"]))
        idx.scan_tables_dev(k, Q, tables, gD[r], gS[r], gL[r])
    idx.set_shard(0, ix["nc"])
    D1 = torch.empty_like(D0); L1 = torch.empty_like(L0)
    idx.merge_partials_dev(k, nq, nshards, gD, gS, gL, D1, L1)
    torch.cuda.synchronize()
    assert torch.equal(L0, L1) and torch.equal(D0, D1)
    idx.close()


@pytest.mark.parametrize("name", ["adc_d32", "grp_d32", "grp_sift_m16", "grp_long_m16", "grp_deep_opq"])
@pytest.mark.parametrize("nprobe,max_codes,k", [(1, 10 ** 9, 1), (3, 1, 5), ("all", 10 ** 9, 10), ("all", 700, 33),
                                                (7, 50, 129)])
def test_operating_point_corners(port, name, nprobe, max_codes, k):
    """Corners of the (nprobe, max_codes, k) space on the GPU's own exact candidate lists: one probe, a max_codes that
    the first list already exceeds, every list probed, k on both sides of the 32 / 128 result-list switches."""
    pkg = load_pkg()
    g, ix, runs = load_golden(name)
    scale = float(ix["centroid_norms"].max())
    idx = make_index(pkg, ix)
    idx.load_index_dict(ix)
    np_ = ix["nc"] if nprobe == "all" else nprobe
    idx.nprobe, idx.max_codes = np_, max_codes
    cd, ci = idx.coarse(g["Q"], np_)
    for pr in ((False, True) if ix["nsubc"] else (False,)):
        idx.do_pruning = pr
        D, L = idx.search2(k, g["Q"], cd, ci)
        Dp, Lp = port.search(ix, g["Q"], cd, ci, k, np_, max_codes, pr)
        o = np.argsort(Dp, axis=1, kind="stable")
        Dps, Lps = np.take_along_axis(Dp, o, 1), np.take_along_axis(Lp, o, 1)
        assert dist_close(D, Dps, scale), (name, pr)
        assert labels_match_up_to_ties(D, L, Dps, Lps, scale) == 0, (name, pr)
    idx.close()


@pytest.mark.parametrize("impl", [1, 2, 3])
@pytest.mark.parametrize("nprobe", [1, 40])
def test_coarse_long_rows_vs_brute_force(nprobe, impl):
    """nc large enough (>= 2^18) for the select kernel's long-row path (threshold from the per-thread minima) and for
    the GEMM epilogue that keeps 3 values per group."""
    pkg = load_pkg()
    d, nc, nq = 32, 300_000, 200
    rng = np.random.default_rng(5)
    C = rng.standard_normal((nc, d)).astype(np.float32)
    Q = (C[rng.integers(0, nc, nq)] + 0.5 * rng.standard_normal((nq, d))).astype(np.float32)
    idx = pkg.IndexIVF_HNSW(d, nc, 8, 8)
    idx.build_quantizer(centroids=C)
    idx.set_option("coarse_impl", impl)
    cd, ci = idx.coarse(Q, nprobe)
    Q64, C64 = Q.astype(np.float64), C.astype(np.float64)
    dd = (Q64 * Q64).sum(1)[:, None] - 2.0 * (Q64 @ C64.T) + (C64 * C64).sum(1)[None, :]
    want = np.sort(dd, axis=1)[:, :nprobe]
    assert np.allclose(cd, want, rtol=2e-5, atol=1e-4)
    got = np.take_along_axis(dd, ci.astype(np.int64), 1)
    assert np.allclose(got, want, rtol=2e-5, atol=1e-4)
    assert np.all(np.diff(cd, axis=1) >= 0)
    idx.close()


@pytest.mark.parametrize("W", [3, 24])
@pytest.mark.parametrize("name", ["grp_sift_m16", "adc_sift_m16", "grp_long_m16"])
def test_packed_records_sharded_search_equals_single_gpu(name, W):
    """The all-to-all form of the sharded search on ONE GPU: records -> one packed stream per owning rank
    (pack_records_dev, checked word for word against the numpy statement tests/pack_ref.py), every 'rank' scans its
    lists from its stream (scan_packed_dev), the merged answer equals the unsharded search bit for bit.  With 24
    'ranks' most (query, rank) pairs hold no probe at all: the scan's early exit for such queries."""
    import torch
    import pack_ref
    pkg = load_pkg()
    sharding = importlib.import_module("ivf-hnsw_b200.sharding")
    g, ix, runs = load_golden(name)
    k, nprobe = int(g["k"]), int(g["nprobe"])
    idx = make_index(pkg, ix)
    idx.load_index_dict(ix)
    idx.nprobe = nprobe
    tag, mc, pr = runs[-1]
    idx.max_codes, idx.do_pruning = mc, pr
    dev = torch.device("cuda:0")
    Q = torch.from_numpy(np.ascontiguousarray(g["Q"])).to(dev)
    nq = Q.shape[0]
    D0 = torch.empty(nq, k, device=dev); L0 = torch.empty(nq, k, dtype=torch.int64, device=dev)
    idx.search_dev(k, Q, D0, L0)
    words = idx.tables_bytes() // 4
    assert words == pack_ref.record_words(nprobe, ix["nsubc"])
    tables = torch.zeros(nq, words, dtype=torch.int32, device=dev)
    idx.prepare_dev(Q, 0, nq, tables)
    packed = torch.full((nq * words,), -1, dtype=torch.int32, device=dev)
    cnt = torch.zeros(W, nq, dtype=torch.int32, device=dev)
    off = torch.zeros(W, nq, dtype=torch.int32, device=dev)
    pw = torch.zeros(W, dtype=torch.int64, device=dev)
    idx.pack_records_dev(nq, tables, W, packed, cnt, off, pw)
    torch.cuda.synchronize()
    streams, rcnt, roff, rpw = pack_ref.pack_records(tables.cpu().numpy().view(np.uint32), nprobe, ix["nsubc"], ix["nc"], W)
    assert np.array_equal(cnt.cpu().numpy().view(np.uint32), rcnt)
    assert np.array_equal(off.cpu().numpy().view(np.uint32), roff)
    assert np.array_equal(pw.cpu().numpy().astype(np.uint64), rpw)
    assert np.array_equal(packed.cpu().numpy().view(np.uint32)[:int(rpw.sum())], np.concatenate(streams))
    # every 'rank' scans from its own stream; here all queries have one home, so the stream IS the receive buffer
    gD = torch.empty(W, nq, k, device=dev); gS = torch.empty(W, nq, k, dtype=torch.int32, device=dev)
    gL = torch.empty(W, nq, k, dtype=torch.int64, device=dev)
    base = np.concatenate([[0], np.cumsum(rpw)[:-1]]).astype(np.int64)
    for r in range(W):
        idx.set_shard(*sharding.shard_range(r, W, ix["nc"]))
        stream = packed[int(base[r]):int(base[r] + rpw[r])].contiguous() if rpw[r] else torch.zeros(1, dtype=torch.int32, device=dev)
        idx.scan_packed_dev(k, Q, stream, off[r].contiguous(), cnt[r].contiguous(), gD[r], gS[r], gL[r])
    D1 = torch.empty_like(D0); L1 = torch.empty_like(L0)
    idx.merge_partials_dev(k, nq, W, gD, gS, gL, D1, L1)
    torch.cuda.synchronize()
    assert torch.equal(L0, L1) and torch.equal(D0, D1)
    idx.close()
