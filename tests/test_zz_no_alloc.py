"""ivfhnsw_b200_coarse_noalloc_dev (no counterpart in the reference):
the exact coarse search with its scratch in the library's static device array must return what the allocating
entry point ivfhnsw_b200_coarse_dev returns -- on two calls in a row that share the array -- and must reject sizes
beyond the array.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from conftest import load_pkg

pytestmark = pytest.mark.gpu


def _both(idx, torch, Q, nprobe):
    dev = Q.device
    nq = Q.shape[0]
    d0 = torch.empty(nq, nprobe, device=dev)
    i0 = torch.empty(nq, nprobe, dtype=torch.int32, device=dev)
    idx.coarse_dev(Q, nprobe, d0, i0)                       # existing entry point (workspace from cudaMalloc)
    # poison the outputs so that a call that wrote nothing cannot pass
    d1 = torch.full((nq, nprobe), -1.0, device=dev)
    i1 = torch.full((nq, nprobe), -1, dtype=torch.int32, device=dev)
    flags = torch.full((1,), 0x7fffffff, dtype=torch.int32, device=dev)
    idx.coarse_noalloc_dev(Q, nprobe, d1, i1, flags)        # static scratch, allocates nothing
    torch.cuda.synchronize()
    assert int(flags.item()) == 0, "select overflow flags %#x" % int(flags.item())
    return d0.cpu().numpy(), i0.cpu().numpy(), d1.cpu().numpy(), i1.cpu().numpy()


def test_coarse_noalloc_equals_coarse_dev_twice():
    import torch
    pkg = load_pkg()
    rng = np.random.default_rng(77)
    d, nc = 64, 5000                                         # ragged: 40 groups, the last one partial
    cent = rng.standard_normal((nc, d), dtype=np.float32) * 3
    idx = pkg.IndexIVF_HNSW(d, nc, 8, 8)
    idx.build_quantizer(centroids=cent)
    dev = torch.device("cuda:0")
    try:
        # two calls in a row with different shapes: the second finds the first one's values in the shared array
        for nq, nprobe, seed in ((700, 32, 1), (333, 7, 2)):
            Q = torch.from_numpy(np.random.default_rng(seed).standard_normal((nq, d), dtype=np.float32) * 3).to(dev)
            d0, i0, d1, i1 = _both(idx, torch, Q, nprobe)
            assert np.array_equal(i0, i1)
            assert np.allclose(d0, d1, rtol=1e-4, atol=0)   # tolerance of the repo's distance checks (north_star)
            assert np.array_equal(d0, d1)                    # in fact the same fp32 re-rank: bit-equal
            # and both equal a float64 brute force
            q = Q.cpu().numpy().astype(np.float64)
            ref = ((q[:, None, :] - cent[None].astype(np.float64)) ** 2).sum(-1)
            assert np.array_equal(np.argsort(ref, axis=1, kind="stable")[:, :nprobe], i1.astype(np.int64))
        # beyond the static array: rejected with a clear message, nothing computed
        nq_big = 8 * 1024 * 1024 // (40 * 32) + 64
        Qb = torch.zeros(nq_big, d, device=dev)
        db = torch.empty(nq_big, 4, device=dev)
        ib = torch.empty(nq_big, 4, dtype=torch.int32, device=dev)
        with pytest.raises(pkg.B200Error, match="scratch"):
            idx.coarse_noalloc_dev(Qb, 4, db, ib)
    finally:
        idx.close()
