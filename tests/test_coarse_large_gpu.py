"""K-A2 at the centroid counts of the benchmarked configurations (2^18, 2^20): the select kernel's long-row path
(threshold from per-thread minima, coarse_select_kernel<16>), the 3-values-per-group GEMM epilogue and the
self-excluding neighbour search of K-G3 -- against a float64 brute force computed with torch on the same GPU."""
import numpy as np
import pytest

from conftest import load_pkg

pytestmark = pytest.mark.gpu


def _brute(torch, Q, C, dk):
    Q64, C64 = Q[:, :dk].double(), C[:, :dk].double()
    return (Q64 * Q64).sum(1)[:, None] - 2.0 * (Q64 @ C64.T) + (C64 * C64).sum(1)[None, :]


@pytest.mark.parametrize("nc,d", [(1 << 18, 32), (1 << 20, 64)])
def test_coarse_at_bench_centroid_counts(nc, d):
    import torch
    pkg = load_pkg()
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev)
    g.manual_seed(nc + d)
    C = torch.randn(nc, d, device=dev, generator=g)
    nq = 300
    Q = C[torch.randint(0, nc, (nq,), device=dev, generator=g)] + 0.5 * torch.randn(nq, d, device=dev, generator=g)
    Q[1] = 0                                              # a zero query next to full-norm ones (one key scale per batch)
    Q[2] = Q[2].round()                                   # exactly representable in fp16: no lo part
    idx = pkg.IndexIVF_HNSW(d, nc, 8, 8)
    idx.build_quantizer(centroids=C.cpu().numpy())
    dd = _brute(torch, Q, C, (d // 16) * 16)
    for impl in (1, 2, 3):
        idx.set_option("coarse_impl", impl)
        # 512 probes at 2^20 centroids: the un-sampled, un-cached radix select over 8 192 keys and a long partner list
        for nprobe in (1, 64, 128) + ((512,) if (impl == 3 and nc == 1 << 20) else ()):
            cd = torch.empty(nq, nprobe, device=dev)
            ci = torch.empty(nq, nprobe, dtype=torch.int32, device=dev)
            idx.coarse_dev(Q, nprobe, cd, ci)
            idx.sync()                                   # also fails on a select overflow flag
            want = torch.sort(dd, dim=1).values[:, :nprobe]
            got = torch.gather(dd, 1, ci.long())
            assert torch.allclose(cd.double(), want, rtol=2e-5, atol=1e-4), (impl, nprobe)
            assert torch.allclose(got, want, rtol=2e-5, atol=1e-4), (impl, nprobe)
            assert bool((cd[:, 1:] >= cd[:, :-1]).all())
            # distinct ids per row
            assert int((torch.sort(ci, dim=1).values.diff(dim=1) == 0).sum()) == 0
    idx.stats()                                          # raises when the sticky overflow flags are set
    idx.close()


def test_neighbour_lists_exclude_self_at_2_18():
    """K-G3 (IndexIVF_HNSW_Grouping.cpp:47-63): the nsubc nearest OTHER centroids of every centroid, nearest first,
    self dropped by id; checked on 2 000 sampled rows of a 2^18-centroid table (with duplicated centroids: the
    duplicate is a legitimate neighbour at distance 0, the centroid itself is not)."""
    import torch
    pkg = load_pkg()
    dev = torch.device("cuda:0")
    nc, d, nsubc = 1 << 18, 32, 16
    g = torch.Generator(device=dev)
    g.manual_seed(99)
    C = torch.randn(nc, d, device=dev, generator=g)
    C[1000:1010] = C[5000:5010]                          # exact duplicates
    idx = pkg.IndexIVF_HNSW_Grouping(d, nc, 8, 8, nsubc)
    idx.build_quantizer(centroids=C.cpu().numpy())
    idx.finalize()                                       # empty index: nn lists and icd only
    out = idx.export_index_dict()
    rows = torch.cat([torch.arange(990, 1020, device=dev), torch.randint(0, nc, (1970,), device=dev, generator=g)])
    dd = _brute(torch, C[rows], C, d)
    dd[torch.arange(rows.numel(), device=dev), rows] = float("inf")      # self
    want = torch.sort(dd, dim=1).values[:, :nsubc]
    nn = torch.from_numpy(out["nn_idx"].astype(np.int64)).to(dev)[rows]
    icd = torch.from_numpy(out["icd"]).to(dev)[rows]
    assert not bool((nn == rows[:, None]).any())
    assert torch.allclose(torch.gather(dd, 1, nn), want, rtol=2e-5, atol=1e-4)
    assert torch.allclose(icd.double(), want, rtol=2e-5, atol=1e-4)
    idx.close()


@pytest.mark.parametrize("exclude_self", [False, True])
def test_adjacent_columns_both_among_the_nearest(exclude_self):
    """impl 3 at >= 2048 groups reduces every column pair (2j, 2j+1) to its smaller key in the GEMM epilogue; the select
    stage must bring the other column back.  Worst case on purpose: centroid 2j+1 is a slightly perturbed copy of
    centroid 2j, so BOTH columns of a pair are among the nearest of every query -- including the pair that holds the
    query's own centroid in the self-excluding search of K-G3."""
    import torch
    pkg = load_pkg()
    dev = torch.device("cuda:0")
    nc, d = 1 << 18, 32
    g = torch.Generator(device=dev)
    g.manual_seed(4242)
    C = torch.randn(nc, d, device=dev, generator=g)
    C[1::2] = C[0::2] + 1e-3 * torch.randn(nc // 2, d, device=dev, generator=g)
    if exclude_self:
        nsubc = 16
        idx = pkg.IndexIVF_HNSW_Grouping(d, nc, 8, 8, nsubc)
        idx.build_quantizer(centroids=C.cpu().numpy())
        idx.finalize()
        out = idx.export_index_dict()
        rows = torch.randint(0, nc, (2000,), device=dev, generator=g)
        dd = _brute(torch, C[rows], C, d)
        dd[torch.arange(rows.numel(), device=dev), rows] = float("inf")
        want = torch.sort(dd, dim=1).values[:, :nsubc]
        nn = torch.from_numpy(out["nn_idx"].astype(np.int64)).to(dev)[rows]
        assert not bool((nn == rows[:, None]).any())
        assert bool((nn[:, 0] == (rows ^ 1)).all())      # the perturbed copy is the nearest OTHER centroid
        assert torch.allclose(torch.gather(dd, 1, nn), want, rtol=2e-5, atol=1e-4)
    else:
        idx = pkg.IndexIVF_HNSW(d, nc, 8, 8)
        idx.build_quantizer(centroids=C.cpu().numpy())
        nq = 400
        Q = C[torch.randint(0, nc, (nq,), device=dev, generator=g)] + 0.3 * torch.randn(nq, d, device=dev, generator=g)
        dd = _brute(torch, Q, C, d)
        for nprobe in (1, 2, 32, 128):
            cd = torch.empty(nq, nprobe, device=dev)
            ci = torch.empty(nq, nprobe, dtype=torch.int32, device=dev)
            idx.coarse_dev(Q, nprobe, cd, ci)
            idx.sync()
            want = torch.sort(dd, dim=1).values[:, :nprobe]
            assert torch.allclose(torch.gather(dd, 1, ci.long()), want, rtol=2e-5, atol=1e-4), nprobe
            assert torch.allclose(cd.double(), want, rtol=2e-5, atol=1e-4), nprobe
            assert int((torch.sort(ci, dim=1).values.diff(dim=1) == 0).sum()) == 0
            if nprobe >= 2:   # both columns of the nearest pair are returned
                assert float(((ci[:, 0] ^ 1) == ci[:, 1]).float().mean()) > 0.9
    idx.stats()
    idx.close()


def test_locality_ordered_centroid_table():
    """A centroid file whose ids follow the geometry (hierarchical k-means output, a space-filling order): the nprobe
    nearest centroids of a query are then ~nprobe CONSECUTIVE ids.  The tensor-core GEMM keeps a few values per group
    of 128 columns, so its columns are a fixed low-discrepancy permutation of the ids (TcSide::idmul): neighbouring ids
    never share a group and nothing overflows.  Worst case here: the centroids are a random WALK in id order."""
    import torch
    pkg = load_pkg()
    dev = torch.device("cuda:0")
    nc, d, nq = 1 << 18, 32, 300
    g = torch.Generator(device=dev)
    g.manual_seed(77)
    # AR(1) walk in id order: centroid i+k is at squared distance ~2 (1 - rho^k) ||c||^2 from centroid i, so the 128
    # nearest of a query sitting on centroid i are ids i-64 .. i+64, two fp16 margins apart from each other
    from scipy.signal import lfilter
    rho = 0.999
    rng = np.random.default_rng(77)
    walk = lfilter([np.sqrt(1 - rho * rho)], [1.0, -rho], rng.standard_normal((nc + 4096, d)), axis=0)[4096:]
    C = torch.from_numpy(np.ascontiguousarray(walk, dtype=np.float32)).to(dev)
    Q = C[torch.randint(0, nc, (nq,), device=dev, generator=g)] + 0.001 * torch.randn(nq, d, device=dev, generator=g)
    dd = _brute(torch, Q, C, d)
    for cls, nc_use in ((pkg.IndexIVF_HNSW, nc), (pkg.IndexIVF_HNSW, 100_000)):     # 100 000: not a power of two
        Cu = C[:nc_use].contiguous()
        idx = cls(d, nc_use, 8, 8)
        idx.build_quantizer(centroids=Cu.cpu().numpy())
        ddu = dd[:, :nc_use]
        for impl in (2, 3):
            idx.set_option("coarse_impl", impl)
            for nprobe in (1, 128):
                cd = torch.empty(nq, nprobe, device=dev)
                ci = torch.empty(nq, nprobe, dtype=torch.int32, device=dev)
                idx.coarse_dev(Q, nprobe, cd, ci)
                idx.sync()                                   # raises on a select overflow flag
                want = torch.sort(ddu, dim=1).values[:, :nprobe]
                assert torch.allclose(torch.gather(ddu, 1, ci.long()), want, rtol=2e-5, atol=1e-4), (nc_use, impl, nprobe)
                assert int((torch.sort(ci, dim=1).values.diff(dim=1) == 0).sum()) == 0
        idx.stats()
        idx.close()
