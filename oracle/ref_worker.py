"""TEST/BENCH INFRASTRUCTURE -- one worker process of the CPU reference arm.

    ref_worker.py <workdir> build            build + save the HNSW over the centroid file (reference code)
    ref_worker.py <workdir> search <rank>    load everything, search this rank's query slice, print one JSON line

Runs the COMPILED REFERENCE (oracle/_ref/libivfhnsw_ref.so): IndexIVF_HNSW[_Grouping]::build_quantizer, read,
rotate_quantizer and the sequential search loop of tests/test_ivfhnsw_sift1b.cpp:192-213.  No GPU code here.
"""
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import oracle as O  # noqa: E402


def main():
    workdir, mode = sys.argv[1], sys.argv[2]
    w = json.load(open(os.path.join(workdir, "cfg.json")))
    p = lambda n: os.path.join(workdir, n).encode()
    r = O.Ref(w["d"], w["nc"], w["M"], w["nsubc"])
    devnull = os.open(os.devnull, os.O_WRONLY)
    saved = os.dup(1)
    os.dup2(devnull, 1)      # the reference prints progress to stdout
    try:
        # loads info/edges when they exist, else constructs and saves them (IndexIVF_HNSW.cpp:40-65)
        r.L.ref_build_quantizer(r.h, p("centroids.fvecs"), p("hnsw.info"), p("hnsw.edges"), 16, 120)
        if mode == "build":
            return
        rank = int(sys.argv[3])
        r.L.ref_load_pq(r.h, p("pq.pq"), p("norm.pq"), p("opq.mat") if w["opq"] else b"")
        r.set_do_opq(w["opq"])
        r.read(os.path.join(workdir, "x.index"))
        if w["opq"]:
            r.rotate_quantizer()
        r.set_params(w["nprobe"], w["max_codes"], w["efs"], w["pruning"])
        Q = np.load(os.path.join(workdir, "queries.npy"))
        gt = np.load(os.path.join(workdir, "gt0.npy"))
        n = Q.shape[0]
        lo, hi = rank * n // w["nproc"], (rank + 1) * n // w["nproc"]
        Qs = np.ascontiguousarray(Q[lo:hi])
        k = w["k"]
        for _ in range(w.get("warmup", 0)):
            r.search(Qs[: max(1, (hi - lo) // 4)], k, want_coarse=False)
        steps = max(1, w.get("steps", 1))
        t0 = time.perf_counter()
        for _ in range(steps):
            D, L = r.search(Qs, k, want_coarse=False)
        sec = time.perf_counter() - t0
        correct = int(sum(gt[lo + i] in L[i] for i in range(hi - lo)))
    finally:
        os.dup2(saved, 1)
    if mode == "search":
        print(json.dumps({"n": (hi - lo) * steps, "sec": sec, "correct": correct * steps}), flush=True)


if __name__ == "__main__":
    main()
