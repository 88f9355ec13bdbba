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
    # the default is configs[1]; the SIFT1B-shape configuration is covered from both sides
    assert bench.WORKLOADS["c2"]["nc"] == 65536 and bench.WORKLOADS["c2"]["nsubc"] == 64
    assert bench.WORKLOADS["c4s"]["nprobe"] == bench.WORKLOADS["c4n"]["nprobe"] == 128
    assert bench.WORKLOADS["c4n"]["nc"] == 1 << 20 and bench.WORKLOADS["c3"]["nc"] == 1 << 20


def test_kmeans_helper_imports_without_a_gpu():
    kmeans = importlib.import_module("ivf-hnsw_b200.kmeans")
    assert callable(kmeans.lloyd_centroids)
