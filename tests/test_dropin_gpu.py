"""The drop-in C++ API (include/ivf-hnsw/*.h + libivfhnsw_dropin.so) driven by a C++ program that is written
like the reference's own mains (tests/cpp/dropin_smoke.cpp): files on disk, train_pq, assign, add_batch /
add_group from an OpenMP loop, write, read into a fresh index, one search() call per query, recall."""
import os
import subprocess
import tempfile

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "ivf-hnsw_b200", "dropin_smoke")


@pytest.mark.gpu
@pytest.mark.parametrize("mode", [("adc",), ("grouping",), ("adc", "opq"), ("grouping", "opq")])
def test_cpp_dropin_end_to_end(mode):
    assert os.path.exists(BIN), "build() did not produce dropin_smoke"
    with tempfile.TemporaryDirectory() as tmp:
        r = subprocess.run([BIN, tmp] + list(mode), capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "dropin_smoke OK" in r.stdout


def test_reference_mains_compile_against_the_dropin_headers():
    """Source compatibility of the boundary: the reference's UNMODIFIED tests/*.cpp compile and link against
    include/ivf-hnsw, include/faiss, include/hnswlib (only where the reference tree is mounted)."""
    if not os.path.isdir("/root/reference/tests"):
        pytest.skip("reference tree not mounted on this machine")
    if not os.path.exists(os.path.join(ROOT, "ivf-hnsw_b200", "libivfhnsw_dropin.so")):
        pytest.skip("drop-in library not built (run __graft_entry__.build())")
    r = subprocess.run(["make", "-C", os.path.join(ROOT, "ivf-hnsw_b200", "csrc"), "refcheck"], capture_output=True,
                       text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("refcheck OK") == 12      # every main of the reference's tests/ directory


def _write_bvecs(path, a):
    import numpy as np
    n, d = a.shape
    out = np.empty((n, 4 + d), np.uint8)
    out[:, :4] = np.frombuffer(np.uint32(d).tobytes(), np.uint8)
    out[:, 4:] = a
    out.tofile(path)


@pytest.mark.gpu
def test_reference_mkindex_main_runs_unmodified():
    """tests/mkindex.cpp of the reference (the "mkindex path" of BASELINE configs[2]), compiled UNMODIFIED against the
    drop-in headers by `make refcheck`, run end to end on a toy data set: train_pq, assign in 1M batches, the bounded
    add_group passes, write, write2 (ORCV directory), rotate_quantizer, dump_inter_centroid_dists."""
    import numpy as np
    import struct
    exe = os.path.join(ROOT, "ivf-hnsw_b200", "csrc", "build", "refcheck", "mkindex")
    if not os.path.exists(exe):
        pytest.skip("refcheck binaries are built where the reference tree is mounted")
    import sys
    sys.path.insert(0, ROOT)
    from conftest import load_pkg
    pkg = load_pkg()
    rng = np.random.default_rng(3)
    d, nc, nb, nt, nsubc = 32, 1024, 1_000_000, 100_000, 16
    cent = rng.integers(20, 235, (nc, d)).astype(np.float32)
    base = np.clip(cent[rng.integers(0, nc, nb)] + rng.normal(0, 9, (nb, d)), 0, 255).astype(np.uint8)
    learn = np.clip(cent[rng.integers(0, nc, nt)] + rng.normal(0, 9, (nt, d)), 0, 255).astype(np.uint8)
    with tempfile.TemporaryDirectory() as tmp:
        p = lambda n: os.path.join(tmp, n)
        _write_bvecs(p("base.bvecs"), base)
        _write_bvecs(p("learn.bvecs"), learn)
        fv = np.empty((nc, d + 1), np.float32)
        fv[:, 0] = np.frombuffer(np.uint32(d).tobytes(), np.float32)[0]
        fv[:, 1:] = cent
        fv.tofile(p("centroids.fvecs"))
        open(p("hnsw.edges"), "wb").write(b"edges")       # write2 copies the edge file
        os.mkdir(p("home_dir"))
        args = [exe, "-M", "16", "-efConstruction", "100", "-nb", str(nb), "-nc", str(nc), "-nsubc", str(nsubc), "-nt", str(nt),
                "-nsubt", "30000", "-d", str(d), "-code_size", "8", "-opq", "on", "-path_base", p("base.bvecs"),
                "-path_learn", p("learn.bvecs"), "-path_centroids", p("centroids.fvecs"), "-path_precomputed_idx",
                p("idx.ivecs"), "-path_info", p("hnsw.info"), "-path_edges", p("hnsw.edges"), "-path_pq", p("pq.pq"),
                "-path_opq_matrix", p("opq.mat"), "-path_norm_pq", p("norm.pq"), "-path_index", p("x.index")]
        r = subprocess.run(args, cwd=tmp, capture_output=True, text=True, timeout=900)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        for f in ("x.index", "pq.pq", "norm.pq", "opq.mat", "idx.ivecs", "centriods.log", "home_dir/hdr.vec",
                  "home_dir/cnorms.vec", "home_dir/pq.vec", "home_dir/normpq.vec", "home_dir/opq.vec", "home_dir/edge.vec"):
            assert os.path.getsize(p(f)) > 0, f
        hdr = struct.unpack("<7I2fB", open(p("home_dir/hdr.vec"), "rb").read()[:37])
        assert hdr[:7] == (nb, nc, 8, 1, d, 16, 100) and hdr[9] == 1
        # the index the main wrote loads back with every vector in it, and its assignment is the exact one
        lab_file = np.fromfile(p("idx.ivecs"), np.uint32)
        assert lab_file[0] == 1_000_000 and lab_file.size == nb + 1
        idx = pkg.IndexIVF_HNSW_Grouping(d, nc, 8, 8, nsubc)
        idx.build_quantizer(centroids=cent)
        idx.read(p("x.index"))
        assert idx.ntotal == nb
        ix = idx.export_index_dict()
        assert np.array_equal(np.diff(ix["list_off"]), np.bincount(lab_file[1:], minlength=nc))
        sample = rng.integers(0, nb, 2000)
        dd = ((base[sample].astype(np.float64)[:, None] - cent.astype(np.float64)[None]) ** 2).sum(-1)
        assert np.mean(dd.argmin(1) == lab_file[1:][sample]) >= 0.999
        idx.close()


@pytest.mark.gpu
def test_rerank_matches_numpy():
    """ivfhnsw_b200_rerank (searchDisk's exact re-rank, IndexIVF_HNSW_Grouping.cpp:365-395): fvec_L2sqr over
    floor(d/16)*16 dims, ascending (distance, label), padding labels last."""
    import numpy as np
    import sys
    sys.path.insert(0, ROOT)
    from conftest import load_pkg
    pkg = load_pkg()
    rng = np.random.default_rng(11)
    d, nq, ncand, k = 40, 37, 20, 10
    idx = pkg.IndexIVF_HNSW(d, 16, 8, 8)
    Q = rng.integers(0, 255, (nq, d)).astype(np.float32)
    cand = rng.integers(0, 255, (nq, ncand, d)).astype(np.float32)
    lab = rng.permutation(nq * ncand).reshape(nq, ncand).astype(np.int64)
    lab[3, 5:] = -1                      # a short result list
    cand[7, 4] = cand[7, 9]              # an exact distance tie: ordered by label
    D, L = idx.rerank(k, Q, cand, lab)
    dk = (d // 16) * 16
    dd = ((Q[:, None, :dk].astype(np.float64) - cand[:, :, :dk].astype(np.float64)) ** 2).sum(-1)
    dd[lab < 0] = np.inf
    for i in range(nq):
        order = sorted(range(ncand), key=lambda j: (dd[i, j], lab[i, j]))[:k]
        want_l = [int(lab[i, j]) if lab[i, j] >= 0 else -1 for j in order]
        assert L[i].tolist() == want_l, i
        fin = np.isfinite(dd[i, order])
        assert np.allclose(D[i][fin], dd[i, order][fin], rtol=1e-6)
    idx.close()
