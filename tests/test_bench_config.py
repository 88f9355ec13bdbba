"""CPU checks of bench.py's workload table (the GPU box only runs it): every workload names a shape the kernels take and
the BASELINE.json configuration it stands for."""
import importlib
import json
import os
import sys

from conftest import ROOT


def test_workloads_are_well_formed():
    sys.path.insert(0, ROOT)
    bench = importlib.import_module("bench")
    base = json.load(open(os.path.join(ROOT, "BASELINE.json")))
    assert len(base["configs"]) == 5
    for name, w in bench.WORKLOADS.items():
        assert w["d"] % w["M"] == 0, name
        assert w["M"] in (8, 16, 32, 64), name
        assert w["d"] % 4 == 0, name
        if w.get("assign"):
            continue
        assert w["nprobe"] <= w["nc"] and w["k"] >= 1, name
        assert w["nprobe"] * max(w["nsubc"], 1) + w["nprobe"] < 65536, name      # uint16 row indices of K-G1
        assert w["nlearn"] >= w["nc"] and w["ntrain"] <= w["nlearn"], name
        assert w["nq"] <= w["nb"], name
    # the default is the configuration the metric is quoted on (configs[3], SIFT1B-shape, 1B x 128, 2^20 centroids)
    c4 = bench.WORKLOADS[bench.DEFAULT_WORKLOAD]
    assert (c4["nb"], c4["nc"], c4["M"], c4["nsubc"], c4["nprobe"], c4["max_codes"], c4["k"]) == \
        (10 ** 9, 1 << 20, 16, 64, 128, 100000, 10) and c4["opq"] and c4["pruning"]
    assert bench.WORKLOADS["c2"]["nc"] == 65536 and bench.WORKLOADS["c2"]["nsubc"] == 64
    assert bench.WORKLOADS["c4s"]["nprobe"] == bench.WORKLOADS["c4n"]["nprobe"] == 128
    assert bench.WORKLOADS["c4n"]["nc"] == 1 << 20 and bench.WORKLOADS["c3"]["nc"] == 1 << 20


def test_kmeans_helper_imports_without_a_gpu():
    kmeans = importlib.import_module("ivf-hnsw_b200.kmeans")
    assert callable(kmeans.lloyd_centroids)


def test_cache_key_depends_on_workload_and_is_stable():
    sys.path.insert(0, ROOT)
    bench = importlib.import_module("bench")
    a = bench.cache_key(bench.WORKLOADS["c4"])
    assert a == bench.cache_key(dict(bench.WORKLOADS["c4"])) and a != bench.cache_key(bench.WORKLOADS["c4s"])
    assert len(bench.csrc_hash()) == 40


def test_generator_is_deterministic_and_ground_truth_is_exact():
    """The synthetic base set is re-drawn batch by batch from seeds (pass 1 / pass 2 of the build, every rank of a
    multi-GPU run): the same (n, seed) must give the same batch; SIFT-shape data is byte-valued, which is what makes
    the harness's TF32/fp32 brute force for the recall figure exact; the streaming ground-truth update equals a float64
    brute force."""
    import numpy as np
    import torch
    sys.path.insert(0, ROOT)
    bench = importlib.import_module("bench")
    w = dict(bench.WORKLOADS["c4s"])
    dev = torch.device("cpu")
    G = bench.Gen(w, torch, dev)
    a, b = G.draw(3000, 99991), G.draw(3000, 99991)
    assert torch.equal(a, b) and not torch.equal(a, G.draw(3000, 99992))
    assert a.dtype == torch.float32 and a.shape == (3000, w["d"])
    assert float(a.min()) >= 0 and float(a.max()) <= 255 and torch.equal(a, a.round())       # byte-valued
    G2 = bench.Gen(w, torch, dev)                      # another process / rank
    assert torch.equal(a, G2.draw(3000, 99991))
    q = G.draw(40, 1235)
    qn = (q * q).sum(1)
    best_d = torch.full((40,), float("inf"))
    best_i = torch.zeros(40, dtype=torch.long)
    chunks = [(0, 3000, 99991), (3000, 2500, 99992 + 3000)]
    allx = []
    for b0, n, seed in chunks:
        x = G.draw(n, seed)
        allx.append(x)
        bench._gt_update(torch, w, q, qn, x, b0, best_d, best_i)
    X = torch.cat(allx).double()
    dd = ((q.double()[:, None, :] - X[None, :, :]) ** 2).sum(-1)
    assert np.array_equal(best_i.numpy(), dd.argmin(1).numpy())
