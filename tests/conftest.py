import importlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
GOLD = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


def load_pkg():
    """The package directory is named ivf-hnsw_b200 (hyphen): import it through importlib."""
    if "ivfhnsw_b200" not in sys.modules:
        sys.modules["ivfhnsw_b200"] = importlib.import_module("ivf-hnsw_b200")
    return sys.modules["ivfhnsw_b200"]


def golden_names():
    return sorted(f[:-4] for f in os.listdir(GOLD) if f.endswith(".npz"))


def load_golden(name):
    """-> (raw npz dict, index dict in the layout oracle.Port and the CUDA library take)."""
    g = dict(np.load(os.path.join(GOLD, name + ".npz")))
    ix = dict(d=int(g["d"]), nc=int(g["nc"]), M=int(g["M"]), nsubc=int(g["nsubc"]),
              centroids=np.ascontiguousarray(g["centroids_search"]),      # search-time table (rotated if OPQ)
              centroids_build=np.ascontiguousarray(g["centroids"]),       # as read from the centroid file
              pqc=g["pqc"], norm_table=g["norm_table"], centroid_norms=g["centroid_norms"],
              list_off=np.ascontiguousarray(g["list_off"], np.int64), ids=g["ids"], codes=g["codes"],
              norm_codes=g["norm_codes"], opq=g["opq"] if int(g["opq_on"]) else None)
    if ix["nsubc"]:
        for k in ("nn_idx", "sub_sizes", "alphas", "icd"):
            ix[k] = np.ascontiguousarray(g[k])
    runs = [(str(t), int(m), bool(p)) for t, m, p in zip(g["runs"], g["runs_max_codes"], g["runs_pruning"])]
    return g, ix, runs


def dist_close(a, b, scale, rtol=1e-4):
    """|a-b| <= rtol * max(|b|, scale): 1e-4 relative (north_star), where `scale` guards the
    catastrophic-cancellation case d ~ 0 (ENN queries): the terms that cancel are O(||c||^2)."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    fin = np.abs(b) < 1e30
    if not np.array_equal(fin, np.abs(a) < 1e30):
        return False
    return bool(np.all(np.abs(a[fin] - b[fin]) <= rtol * np.maximum(np.abs(b[fin]), scale * 1e-2)))


def labels_match_up_to_ties(D, L, Dref, Lref, scale, rtol=1e-4):
    """Label multisets equal after removing elements whose distance ties with the k-th."""
    bad = 0
    for i in range(D.shape[0]):
        if set(L[i].tolist()) == set(Lref[i].tolist()):
            continue
        kth = np.sort(Dref[i])[-1]
        tol = rtol * max(abs(kth), scale * 1e-2)
        a = {l for l, d in zip(L[i], D[i]) if d < kth - tol}
        b = {l for l, d in zip(Lref[i], Dref[i]) if d < kth - tol}
        if a != b:
            bad += 1
    return bad


@pytest.fixture(scope="session")
def port():
    import oracle as O
    O.build(port=True, ref=True)
    return O.Port()
