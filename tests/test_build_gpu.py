"""GPU parity tests of the indexing path (assign, add_batch, add_group, train_pq, write/read)."""
import os
import tempfile

import numpy as np
import pytest

from conftest import golden_names, load_golden, load_pkg
from test_oracle import regenerate_base
from test_search_gpu import make_index

import oracle as O

pytestmark = pytest.mark.gpu


def l2(a, b):
    return ((a - b) ** 2).sum(-1)


@pytest.mark.parametrize("name", golden_names())
def test_assign_exact(port, name):
    """K-A1: exact argmin.  Identical to the port's brute force and never farther than what the
    reference's HNSW found (||x - c_ours||^2 <= ||x - c_ref||^2 (1 + 1e-6))."""
    pkg = load_pkg()
    g, ix, _ = load_golden(name)
    idx = make_index(pkg, ix, search_time=False)
    X = g["X_sub"]
    lab = idx.assign(X)
    plab = port.assign(X, ix["centroids_build"])
    C = ix["centroids_build"].astype(np.float64)
    d_ours = l2(X.astype(np.float64), C[lab])
    d_port = l2(X.astype(np.float64), C[plab])
    d_ref = l2(X.astype(np.float64), C[g["assign_ref"]])
    assert np.mean(lab == plab) >= 0.995
    assert np.all(d_ours <= d_port * (1 + 1e-5) + 1e-6)
    assert np.all(d_ours <= d_ref * (1 + 1e-6) + 1e-6)
    if name.startswith("adc_sift"):   # byte input path
        lab8 = idx.assign(X.astype(np.uint8))
        assert np.array_equal(lab8, lab)
    # k > 1: the k-th nearest (IndexIVF_HNSW.cpp:71 takes the top of a k-element max-heap)
    lab3 = idx.assign(X[:500], k=3)
    dd = l2(X[:500].astype(np.float64)[:, None, :], C[None, :, :]) if C.shape[0] <= 4096 else None
    if dd is not None:
        third = np.sort(dd, axis=1)[:, 2]
        assert np.allclose(l2(X[:500].astype(np.float64), C[lab3]), third, rtol=1e-6, atol=1e-6)
    idx.close()


@pytest.mark.parametrize("name", [n for n in golden_names() if not n.startswith("grp")])
def test_add_batch_matches_reference_lists(name):
    """add_batch with the reference's own assignment and codebooks: same list sizes, same id order
    inside every list (stable scatter), codes byte-identical except fp near-ties."""
    pkg = load_pkg()
    g, ix, _ = load_golden(name)
    base = regenerate_base(name, g)
    idx = make_index(pkg, ix, search_time=False)
    n = base.shape[0]
    half = n // 2   # two batches: exercises incremental staging
    ids = np.arange(n, dtype=np.uint32)
    idx.add_batch(base[:half], ids[:half], g["assign_all"][:half])
    idx.add_batch(base[half:], ids[half:], g["assign_all"][half:])
    idx.compute_centroid_norms()
    out = idx.export_index_dict()
    assert np.array_equal(out["list_off"], ix["list_off"])
    assert np.array_equal(out["ids"], ix["ids"])
    same = np.mean(np.all(out["codes"] == ix["codes"], axis=1))
    assert same >= 0.97, same
    assert np.mean(out["norm_codes"] == ix["norm_codes"]) >= 0.96
    assert np.allclose(out["centroid_norms"], ix["centroid_norms"], rtol=1e-6)
    # reconstruction error of our codes is not worse than the reference's
    pqc = ix["pqc"].reshape(ix["M"], 256, -1)

    def recon_err(codes):
        rec = np.concatenate([pqc[m][codes[:, m]] for m in range(ix["M"])], axis=1)
        order = ix["ids"]
        res = base[order] - ix["centroids_build"][np.repeat(np.arange(ix["nc"]), np.diff(ix["list_off"]))]
        if ix["opq"] is not None:
            res = res @ ix["opq"].T
        return float(((res - rec) ** 2).sum())

    assert recon_err(out["codes"]) <= recon_err(ix["codes"]) * (1 + 1e-5)
    idx.close()


@pytest.mark.parametrize("name", [n for n in golden_names() if n.startswith("grp")])
def test_add_group_matches_reference(name):
    """add_group for every group with the reference's assignment: nn lists (where the HNSW was exact),
    alphas, sub-group sizes, id order and codes."""
    pkg = load_pkg()
    g, ix, _ = load_golden(name)
    base = regenerate_base(name, g)
    idx = make_index(pkg, ix, search_time=False)
    assign = g["assign_all"]
    ids = np.arange(base.shape[0], dtype=np.uint32)
    for grp in range(ix["nc"]):
        m = assign == grp
        idx.add_group(grp, base[m], ids[m])
    idx.compute_centroid_norms()
    idx.compute_inter_centroid_dists()
    out = idx.export_index_dict()
    assert np.array_equal(out["list_off"], ix["list_off"])
    nn_same = np.all(out["nn_idx"] == ix["nn_idx"], axis=1)
    assert nn_same.mean() >= 0.9, nn_same.mean()     # HNSW neighbour lists are approximate
    sizes = np.diff(ix["list_off"])
    ok = nn_same & (sizes > 0)
    assert np.allclose(out["icd"][nn_same], ix["icd"][nn_same], rtol=1e-5)
    # one point whose two best sub-centroid directions are an fp near-tie moves a whole group's alpha
    a_same = np.isclose(out["alphas"], ix["alphas"], rtol=1e-4, atol=1e-6)
    assert a_same[ok].mean() >= 0.95, a_same[ok].mean()
    ok &= a_same
    sub_same = np.all(out["sub_sizes"][ok] == ix["sub_sizes"][ok], axis=1)
    assert sub_same.mean() >= 0.97, sub_same.mean()
    good = np.nonzero(ok)[0][sub_same]
    rows = np.concatenate([np.arange(ix["list_off"][c], ix["list_off"][c + 1]) for c in good])
    assert np.mean(out["ids"][rows] == ix["ids"][rows]) >= 0.995
    assert np.mean(np.all(out["codes"][rows] == ix["codes"][rows], axis=1)) >= 0.95
    idx.close()


@pytest.mark.parametrize("name", ["adc_d32", "grp_d32", "grp_deep_opq"])
def test_write_is_readable_by_the_reference_and_back(name):
    """Byte compatibility of write()/read() with IndexIVF_HNSW::read/write: the reference loads the
    file we wrote and returns the golden answers; we load it back and search identically."""
    pkg = load_pkg()
    g, ix, runs = load_golden(name)
    k, nprobe = int(g["k"]), int(g["nprobe"])
    idx = make_index(pkg, ix)
    idx.load_index_dict(ix)
    tmp = tempfile.mkdtemp(prefix="ivfb200_")
    path = os.path.join(tmp, "x.index")
    idx.write(path)
    idx2 = make_index(pkg, ix)
    idx2.read(path)
    tag, mc, pr = runs[-1]
    for i in (idx, idx2):
        i.nprobe, i.max_codes, i.do_pruning = nprobe, mc, pr
    D1, L1 = idx.search2(k, g["Q"], g["cd_" + tag], g["ci_" + tag])
    D2, L2 = idx2.search2(k, g["Q"], g["cd_" + tag], g["ci_" + tag])
    assert np.array_equal(L1, L2) and np.array_equal(D1, D2)
    if O.have_ref():
        r = O.Ref(ix["d"], ix["nc"], ix["M"], ix["nsubc"])
        r.build_quantizer(ix["centroids"], M=16, efConstruction=200)   # search-time (rotated) centroids
        r.load_pq(ix["pqc"], ix["norm_table"], ix["opq"])
        r.set_do_opq(ix["opq"] is not None)
        r.read(path)
        r.set_params(nprobe, mc, 10, pr)
        if ix["nsubc"] == 0:
            Dr, Lr = r.search2(g["Q"], k, g["cd_" + tag], g["ci_" + tag])
            assert np.array_equal(Lr, g["L_" + tag]) and np.allclose(Dr, g["D_" + tag], rtol=1e-6)
        else:   # Grouping has no search2: check the loaded structures instead
            ex = r.export()
            for key in ("ids", "codes", "norm_codes", "nn_idx", "sub_sizes", "alphas", "icd", "centroid_norms"):
                assert np.array_equal(ex[key], ix[key]), key
        r.close()
    idx.close()
    idx2.close()


@pytest.mark.parametrize("name", ["adc_d32", "adc_opq_d32", "grp_d32"])
def test_train_pq_quality(name):
    """train_pq on the GPU (k-means [+ OPQ]) -- statistical parity: quantisation error of the residual PQ
    within 5 % of the reference-trained codebook on the same learn set, norm codebook sorted-compatible."""
    import make_golden
    import synth_np
    pkg = load_pkg()
    g, ix, _ = load_golden(name)
    case = [c for c in make_golden.CASES if c[0] == name][0]
    _, kind, d, nc, M, nsubc, opq, nb, nt, nq, nprobe, k, efs = case
    base, queries, learn, centroids = synth_np.dataset(kind, nb, nq, nt, nc, d, int(g["seed"]))
    idx = make_index(pkg, ix, search_time=False)
    idx.do_opq = opq
    idx.train_pq(learn)
    assert np.all(np.isfinite(idx.pq_centroids)) and np.all(np.isfinite(idx.norm_centroids))
    if opq:
        A = idx.opq_matrix
        assert np.allclose(A @ A.T, np.eye(d), atol=1e-4)     # orthonormal rotation
    # build with our codebooks and with the reference's, compare ADC recall on the fixture's queries
    ids = np.arange(nb, dtype=np.uint32)

    def build_and_search(use_ours):
        j = make_index(pkg, ix, search_time=False)
        if use_ours:
            j.set_pq(idx.pq_centroids, idx.norm_centroids, idx.opq_matrix)
        j.do_opq = opq
        if nsubc:
            for grp in range(nc):
                m = g["assign_all"] == grp
                j.add_group(grp, base[m], ids[m])
        else:
            j.add_batch(base, ids, g["assign_all"])
        j.compute_centroid_norms()
        j.finalize()
        if opq:
            j.rotate_quantizer()
        j.nprobe, j.max_codes, j.do_pruning = nprobe, 10 ** 9, False
        D, L = j.search(10, queries)
        j.close()
        return L

    gt = np.argmin(((queries[:, None, :].astype(np.float64) - base[None].astype(np.float64)) ** 2).sum(-1), 1)
    rec_ours = np.mean([gt[i] in build_and_search(True)[i] for i in range(nq)])
    rec_ref = np.mean([gt[i] in build_and_search(False)[i] for i in range(nq)])
    assert rec_ours >= rec_ref - 0.15, (rec_ours, rec_ref)   # 40 queries: 0.15 = 6 queries of noise
    idx.close()


@pytest.mark.parametrize("name", ["grp_d32", "grp_sift_m16"])
def test_add_group_equals_port_strict(port, name):
    """K-G2 against the C restatement (fixed fp32 order, no -Ofast), fed the GPU's own exact neighbour lists:
    alpha to 1e-6, sub-group sizes and id order EXACTLY, for every group -- the pin of compute_alpha /
    compute_subcentroid_idxs (IndexIVF_HNSW_Grouping.cpp:673-728) on the CUDA path."""
    pkg = load_pkg()
    g, ix, _ = load_golden(name)
    base = regenerate_base(name, g)
    idx = make_index(pkg, ix, search_time=False)
    assign = g["assign_all"]
    ids = np.arange(base.shape[0], dtype=np.uint32)
    for grp in range(ix["nc"]):
        m = assign == grp
        idx.add_group(grp, base[m], ids[m])
    idx.compute_centroid_norms()
    out = idx.export_index_dict()
    ixb = dict(ix, centroids=ix["centroids_build"])
    code_same, code_n, tested = 0, 0, 0
    for grp in range(ix["nc"]):
        members = np.nonzero(assign == grp)[0]
        if members.size == 0:
            continue
        alpha, sub_idx, oid, oc, on, ss = port.add_group(ixb, grp, out["nn_idx"][grp], out["icd"][grp], base[members], members)
        a, b = out["list_off"][grp], out["list_off"][grp + 1]
        assert abs(alpha - out["alphas"][grp]) <= 1e-6 * max(1.0, abs(alpha)), (grp, alpha, out["alphas"][grp])
        assert np.array_equal(ss, out["sub_sizes"][grp]), grp
        assert np.array_equal(oid, out["ids"][a:b]), grp
        code_same += int(np.all(oc == out["codes"][a:b], axis=1).sum())
        code_n += b - a
        tested += 1
    assert tested >= ix["nc"] // 2
    assert code_same >= 0.99 * code_n, (code_same, code_n)
    idx.close()


def test_streamed_grouping_ingest_equals_one_shot():
    """The bounded-pass ingest (tests/test_ivfhnsw_grouping_sift1b.cpp:138-201): encoding complete groups range by
    range (flush_groups) or through add_group's automatic flush gives the SAME index, bit for bit, as staging
    everything until finalize -- and a group cannot be added twice."""
    import torch
    pkg = load_pkg()
    rng = np.random.default_rng(5)
    d, nc, M, nsubc, n = 32, 96, 8, 16, 30000
    cent = rng.standard_normal((nc, d)).astype(np.float32) * 4
    X = (cent[rng.integers(0, nc, n)] + rng.standard_normal((n, d))).astype(np.float32)
    dev = torch.device("cuda:0")

    def fresh():
        j = pkg.IndexIVF_HNSW_Grouping(d, nc, M, 8, nsubc)
        j.build_quantizer(centroids=cent)
        return j

    a = fresh()
    a.train_pq(X[:8000])
    lab = a.assign(X)
    ids = np.arange(n, dtype=np.uint32)
    Xd, idd, labd = torch.from_numpy(X).to(dev), torch.from_numpy(ids.view(np.int32)).to(dev), torch.from_numpy(lab.view(np.int32)).to(dev)
    a.add_points_dev(Xd, idd, labd)                      # one shot: everything staged until finalize
    a.compute_centroid_norms()
    ref = a.export_index_dict()

    b = fresh()                                          # three passes over group ranges, each flushed
    b.set_pq(a.pq_centroids, a.norm_centroids, None)
    for lo, hi in ((0, 30), (30, 31), (31, nc)):
        m = (labd >= lo) & (labd < hi)
        for part in (m & (idd < n // 2), m & (idd >= n // 2)):     # a range may arrive in several calls
            b.add_points_dev(Xd[part].contiguous(), idd[part].contiguous(), labd[part].contiguous())
        b.flush_groups()
    b.compute_centroid_norms()
    out = b.export_index_dict()

    c = fresh()                                          # add_group with a tiny staging budget: flushes on its own
    c.set_pq(a.pq_centroids, a.norm_centroids, None)
    c.set_option("stage_bytes", 200 * d * 4)
    for grp in range(nc):
        m = lab == grp
        c.add_group(grp, X[m], ids[m])
    c.compute_centroid_norms()
    out2 = c.export_index_dict()
    for o in (out, out2):
        for key in ("list_off", "sub_sizes", "alphas", "nn_idx", "icd", "norm_codes", "codes"):
            assert np.array_equal(o[key], ref[key]), key
        assert np.array_equal(o["ids"], ref["ids"])      # order of arrival inside a sub-group is kept
    # a group that has been encoded cannot receive more points
    e = fresh()
    e.set_pq(a.pq_centroids, a.norm_centroids, None)
    m = lab == 3
    e.add_points_dev(Xd[torch.from_numpy(m).to(dev)].contiguous(), idd[torch.from_numpy(m).to(dev)].contiguous(),
                     labd[torch.from_numpy(m).to(dev)].contiguous())
    e.flush_groups()
    e.add_group(3, X[m][:5], ids[m][:5])
    with pytest.raises(pkg.B200Error, match="COMPLETE groups"):
        e.flush_groups()
    for j in (a, b, c, e):
        j.close()


@pytest.mark.gpu
@pytest.mark.parametrize("grouping", [True, False])
def test_sharded_construction_equals_one_index(grouping):
    """SURVEY.md 8(e), construction side: W 'ranks' (W handles on one GPU here) each set_shard() their range,
    ingest ONLY the points of their own lists, finalize, sum the per-list tables (the all-reduce of
    sharding.exchange_list_meta) and install them; the packed sharded search over those W partial indexes is
    bit-identical to the search of one index that ingested everything.  Also: a complete lists file loaded into a
    sharded handle keeps the owned lists only and answers the same."""
    import torch
    pkg = load_pkg()
    from conftest import ROOT  # noqa: F401
    import importlib
    sharding = importlib.import_module("ivf-hnsw_b200.sharding")
    rng = np.random.default_rng(11)
    d, nc, M, nsubc, n, W, k, nprobe = 32, 120, 8, 8, 40000, 3, 10, 16
    cent = rng.standard_normal((nc, d)).astype(np.float32) * 4
    X = (cent[rng.integers(0, nc, n)] + rng.standard_normal((n, d))).astype(np.float32)
    Q = (cent[rng.integers(0, nc, 300)] + rng.standard_normal((300, d))).astype(np.float32)
    dev = torch.device("cuda:0")

    def fresh():
        j = pkg.IndexIVF_HNSW_Grouping(d, nc, M, 8, nsubc) if grouping else pkg.IndexIVF_HNSW(d, nc, M, 8)
        j.build_quantizer(centroids=cent)
        return j

    full = fresh()
    full.train_pq(X[:8000])
    lab = full.assign(X)
    ids = np.arange(n, dtype=np.uint32)
    Xd = torch.from_numpy(X).to(dev)
    idd = torch.from_numpy(ids.view(np.int32)).to(dev)
    labd = torch.from_numpy(lab.view(np.int32)).to(dev)

    def ingest(j, mask=None):
        if grouping:
            if mask is None:
                j.add_points_dev(Xd, idd, labd)
            else:
                j.add_points_dev(Xd[mask].contiguous(), idd[mask].contiguous(), labd[mask].contiguous())
        else:
            m = np.ones(n, bool) if mask is None else mask.cpu().numpy()
            j.add_batch(X[m], ids[m], lab[m])
        j.compute_centroid_norms()

    ingest(full)
    for j in (full,):
        j.nprobe, j.max_codes = nprobe, 3000
        if grouping:
            j.do_pruning = True
    Qd = torch.from_numpy(Q).to(dev)
    nq = Q.shape[0]
    D0 = torch.empty(nq, k, device=dev); L0 = torch.empty(nq, k, dtype=torch.int64, device=dev)
    full.search_dev(k, Qd, D0, L0)
    torch.cuda.synchronize()
    ref = full.export_index_dict()

    def packed_search(ranks):
        words = ranks[0].tables_bytes() // 4
        tables = torch.zeros(nq, words, dtype=torch.int32, device=dev)
        ranks[0].prepare_dev(Qd, 0, nq, tables)          # any rank can be the home: the metadata is replicated
        packed = torch.zeros(nq * words, dtype=torch.int32, device=dev)
        cnt = torch.zeros(W, nq, dtype=torch.int32, device=dev)
        off = torch.zeros(W, nq, dtype=torch.int32, device=dev)
        pw = torch.zeros(W, dtype=torch.int64, device=dev)
        ranks[0].pack_records_dev(nq, tables, W, packed, cnt, off, pw)
        torch.cuda.synchronize()
        base = np.concatenate([[0], np.cumsum(pw.cpu().numpy())[:-1]]).astype(np.int64)
        gD = torch.empty(W, nq, k, device=dev); gS = torch.empty(W, nq, k, dtype=torch.int32, device=dev)
        gL = torch.empty(W, nq, k, dtype=torch.int64, device=dev)
        for r, j in enumerate(ranks):
            nw = int(pw[r])
            stream = packed[int(base[r]):int(base[r]) + nw].contiguous() if nw else torch.zeros(1, dtype=torch.int32, device=dev)
            j.scan_packed_dev(k, Qd, stream, off[r].contiguous(), cnt[r].contiguous(), gD[r], gS[r], gL[r])
            j.sync()
        D1 = torch.empty_like(D0); L1 = torch.empty_like(L0)
        ranks[0].merge_partials_dev(k, nq, W, gD, gS, gL, D1, L1)
        torch.cuda.synchronize()
        return D1, L1

    # (1) sharded ingest + metadata exchange
    ranks = []
    for r in range(W):
        j = fresh()
        j.set_pq(full.pq_centroids, full.norm_centroids, None)
        lo, hi = sharding.shard_range(r, W, nc)
        j.set_shard(lo, hi)
        ingest(j, (labd >= lo) & (labd < hi))
        j.nprobe, j.max_codes = nprobe, 3000
        if grouping:
            j.do_pruning = True
        ranks.append(j)
    ls = [torch.zeros(nc, dtype=torch.int32, device=dev) for _ in range(W)]
    ss = [torch.zeros(max(nc * nsubc, 1), dtype=torch.int32, device=dev) for _ in range(W)]
    al = [torch.zeros(nc, dtype=torch.float32, device=dev) for _ in range(W)]
    for r, j in enumerate(ranks):
        j.get_list_meta_dev(ls[r], ss[r] if grouping else None, al[r] if grouping else None)
        j.sync()
    LS, SS, AL = sum(ls), sum(ss), sum(al)               # what the all-reduce leaves on every rank
    assert np.array_equal(LS.cpu().numpy(), np.diff(ref["list_off"]).astype(np.int32))
    for j in ranks:
        j.set_list_meta_dev(LS, SS if grouping else None, AL if grouping else None)
        assert j.ntotal == n
    D1, L1 = packed_search(ranks)
    assert torch.equal(L0, L1) and torch.equal(D0, D1)
    # a sharded handle holds its own lists only: no export, no widening of the shard
    with pytest.raises(pkg.B200Error, match="resident"):
        ranks[1].export_index_dict()
    with pytest.raises(pkg.B200Error, match="resident"):
        ranks[1].set_shard(0, nc)
    bad = LS.clone(); lo, hi = sharding.shard_range(1, W, nc); bad[lo] += 1
    with pytest.raises(pkg.B200Error, match="resident list"):
        ranks[1].set_list_meta_dev(bad, SS if grouping else None, AL if grouping else None)
    # (2) a complete index loaded into sharded handles: every rank keeps 1/W of the codes
    loaded = []
    for r in range(W):
        j = fresh()
        j.set_pq(full.pq_centroids, full.norm_centroids, None)
        j.set_shard(*sharding.shard_range(r, W, nc))
        j.load_index_dict(ref)
        j.compute_centroid_norms()
        j.nprobe, j.max_codes = nprobe, 3000
        if grouping:
            j.do_pruning = True
        assert j.ntotal == n
        loaded.append(j)
    D2, L2 = packed_search(loaded)
    assert torch.equal(L0, L2) and torch.equal(D0, D2)
    for j in [full] + ranks + loaded:
        j.close()
