"""CPU tests: the plain-C port (oracle/ivfhnsw_port.c) against the golden fixtures produced by the
COMPILED REFERENCE (oracle/make_golden.py), and against the live reference library where it exists."""
import numpy as np
import pytest

from conftest import golden_names, load_golden, dist_close

import oracle as O


def regenerate_base(name, g):
    """The fixtures keep only the first vectors of the base set; the rest is regenerated from the
    seed with the committed generator and checked against the stored prefix."""
    import make_golden
    import synth_np
    case = [c for c in make_golden.CASES if c[0] == name][0]
    _, kind, d, nc, M, nsubc, opq, nb, nt, nq, nprobe, k, efs = case
    base = synth_np.dataset(kind, nb, nq, nt, nc, d, int(g["seed"]))[0]
    assert np.array_equal(base[:g["X_sub"].shape[0]], g["X_sub"]), "generator drifted from the fixture"
    return base


@pytest.mark.parametrize("name", golden_names())
def test_port_search_matches_reference_heaps(port, name):
    """search / search2 / Grouping::search incl. max_codes cut-off and pruning quirks: the port must
    reproduce the reference's heap contents on the reference's own HNSW candidate lists."""
    g, ix, runs = load_golden(name)
    k, nprobe = int(g["k"]), int(g["nprobe"])
    scale = float(ix["centroid_norms"].max())
    for tag, mc, pr in runs:
        D, L = port.search(ix, g["Q"], g["cd_" + tag], g["ci_" + tag], k, nprobe, mc, pr)
        Dr, Lr = g["D_" + tag], g["L_" + tag]
        assert np.array_equal(L, Lr), (name, tag)          # same labels in the same heap slots
        assert dist_close(D, Dr, scale), (name, tag)


@pytest.mark.parametrize("name", golden_names())
def test_port_exact_coarse_vs_hnsw(port, name):
    """The exact coarse stage must contain everything the HNSW found at least as close: for every
    query, exact dist[j] <= hnsw dist[j] rank by rank (HNSW is approximate), and where the HNSW was
    exact the lists agree."""
    g, ix, runs = load_golden(name)
    nprobe = int(g["nprobe"])
    tag = runs[0][0]
    Q = g["Q"]
    if ix["opq"] is not None:
        Q = port.opq_apply(ix["opq"], Q)
    cd, ci = port.coarse(Q, ix["centroids"], nprobe)
    hd, hi = g["cd_" + tag], g["ci_" + tag]
    assert np.all(cd <= hd * (1 + 1e-5) + 1e-6)
    same = np.mean(ci == hi)
    assert same > 0.9, same
    # identical ids => identical distances up to fp noise
    m = ci == hi
    assert np.allclose(cd[m], hd[m], rtol=1e-5, atol=1e-3)


@pytest.mark.parametrize("name", golden_names())
def test_port_centroid_norms_and_icd(port, name):
    g, ix, _ = load_golden(name)
    cn = port.centroid_norms(ix["centroids_build"])
    assert np.allclose(cn, ix["centroid_norms"], rtol=1e-6)
    if ix["nsubc"]:
        icd = port.inter_centroid_dists(ix["centroids_build"], ix["nn_idx"])
        assert np.allclose(icd, ix["icd"], rtol=1e-5)


@pytest.mark.parametrize("name", [n for n in golden_names() if not n.startswith("grp")])
def test_port_add_batch_codes(port, name):
    """add_batch restatement: codes / norm codes of the first vectors vs what the reference stored."""
    g, ix, _ = load_golden(name)
    X, idx = g["X_sub"], g["assign_ref"]
    ixb = dict(ix, centroids=ix["centroids_build"])
    codes, ncodes, _ = port.add_batch_encode(ixb, X, idx)
    # locate the reference's record of vector i: id == i inside list idx[i]
    off, ids = ix["list_off"], ix["ids"]
    same_c = same_n = 0
    for i in range(X.shape[0]):
        a, b = off[idx[i]], off[idx[i] + 1]
        j = a + int(np.nonzero(ids[a:b] == i)[0][0])
        same_c += int(np.array_equal(ix["codes"][j], codes[i]))
        same_n += int(ix["norm_codes"][j] == ncodes[i])
    # dsub >= 16 goes through BLAS in the reference (different rounding) -> near-ties may flip
    assert same_c >= 0.98 * X.shape[0], same_c
    assert same_n >= 0.97 * X.shape[0], same_n


@pytest.mark.parametrize("name", [n for n in golden_names() if n.startswith("grp")])
def test_port_add_group(port, name):
    """add_group restatement vs the reference: alpha, sub-group sizes, ids order and codes of whole groups."""
    g, ix, _ = load_golden(name)
    ixb = dict(ix, centroids=ix["centroids_build"])
    assign = g["assign_all"]
    base = regenerate_base(name, g)
    tested = 0
    for grp in range(0, ix["nc"], 7):
        members = np.nonzero(assign == grp)[0]
        if members.size == 0:
            continue
        nn, cvn = port.centroid_neighbours(ix["centroids_build"], ix["nsubc"], grp)
        # the HNSW neighbour list of the reference is approximate; only compare when it was exact
        if not np.array_equal(nn, ix["nn_idx"][grp]):
            nn, cvn = ix["nn_idx"][grp], ix["icd"][grp]
        alpha, sub_idx, oid, oc, on, ss = port.add_group(ixb, grp, nn, cvn, base[members], members)
        a, b = ix["list_off"][grp], ix["list_off"][grp + 1]
        assert abs(alpha - ix["alphas"][grp]) <= 1e-5 * max(1.0, abs(alpha)), (grp, alpha, ix["alphas"][grp])
        assert np.array_equal(ss, ix["sub_sizes"][grp]), grp
        assert np.array_equal(oid, ix["ids"][a:b]), grp
        assert np.mean(np.all(oc == ix["codes"][a:b], axis=1)) >= 0.95
        tested += 1
    assert tested >= 1


@pytest.mark.skipif(not O.have_ref(), reason="compiled reference not present")
def test_port_vs_live_reference_small():
    """Fresh data through the live reference and the port (not just the stored fixtures)."""
    import synth_np
    d, nc, M, nq, nprobe, k = 32, 32, 8, 16, 6, 5
    base, queries, learn, centroids = synth_np.dataset("sift", 1500, nq, 1500, nc, d, 99)
    r = O.Ref(d, nc, M, 0)
    r.build_quantizer(centroids)
    r.set_params(nprobe, 10 ** 9, 64, False)
    r.train_pq(learn)
    idx = r.assign(base)
    r.add_batch(base, np.arange(base.shape[0], dtype=np.uint32), idx)
    r.compute_centroid_norms()
    ix = r.export()
    D, L, cd, ci = r.search(queries, k)
    r.close()
    P = O.Port()
    D2, L2 = P.search(ix, queries, cd, ci, k, nprobe, 10 ** 9)
    assert np.array_equal(L, L2)
    assert np.allclose(D, D2, rtol=1e-5)
    # l2sqr restatement vs the reference's AVX kernel
    rng = np.random.default_rng(0)
    R = O.Ref(32, 4, 8)
    for _ in range(50):
        x = (rng.standard_normal(100) * 30).astype(np.float32)
        y = (rng.standard_normal(100) * 30).astype(np.float32)
        a = R.L.ref_fvec_L2sqr(x, y, 100)
        b = P.l2sqr16(x, y)
        assert abs(a - b) <= 1e-6 * abs(a)
    R.close()
