#!/usr/bin/env python
"""BASELINE configs[4]: the ENN variant (examples/run_sift1b_enn_64bits.sh, tests/test_ivfhnsw_sift1b_enn.cpp:27-47:
IVFADC, nprobe 1, k 1, queries = base vectors) swept over the query batch 1 .. 100k, with 64-BYTE codes (what the
script's -code_size 64 means) and 8-byte = 64-BIT codes (what its name says), at 1/50 scale (20M x 128, 2^15 centroids,
list length of SIFT1B kept).  Per batch size: latency and QPS of the C-ABI call with HOST buffers (ivfhnsw_b200_search:
H2D + search + D2H) and of the device-resident call, plus the one-query search() of the C++ drop-in when the smoke
binary is there.  Writes profiles/r02_bench_c5_sweep.json.

    python profiles/c5_sweep.py [--workloads c5s,c5b] [--c4s]"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def sweep(name, torch, dev, batches, reps_of):
    w = dict(bench.WORKLOADS[name])
    log = lambda m: print("[c5 %s] %s" % (name, m), file=sys.stderr, flush=True)
    idx, queries, gt0, _ = bench.build_index(w, torch, dev, log)
    k = w["k"]
    G = bench.Gen(w, torch, dev)
    pool = G.draw(1_000_000, 99991)[:100_000].contiguous() if w.get("enn") else G.draw(100_000, 4242)
    rows = []
    for nq in batches:
        q_dev = pool[:nq].contiguous()
        q_h = q_dev.cpu().pin_memory()
        D = torch.empty(nq, k, device=dev); L = torch.empty(nq, k, dtype=torch.int64, device=dev)
        Dh = torch.empty(nq, k).pin_memory(); Lh = torch.empty(nq, k, dtype=torch.int64).pin_memory()
        reps = reps_of(nq)
        for _ in range(3):
            idx.search_host_buffers(k, q_h, Dh, Lh)
        t0 = time.perf_counter()
        for _ in range(reps):
            idx.search_host_buffers(k, q_h, Dh, Lh)
        host_s = (time.perf_counter() - t0) / reps
        for _ in range(3):
            idx.search_dev(k, q_dev, D, L)
        idx.sync()
        t0 = time.perf_counter()
        for _ in range(reps):
            idx.search_dev(k, q_dev, D, L)
        idx.sync()
        dev_s = (time.perf_counter() - t0) / reps
        hit = float((Lh[:, 0] == torch.arange(nq)).float().mean()) if w.get("enn") else None
        rows.append({"nq": nq, "host_call_us": host_s * 1e6, "host_call_qps": nq / host_s, "device_call_us": dev_s * 1e6,
                     "device_call_qps": nq / dev_s, "recall_at_1": hit})
        log("nq %6d: host %.1f us (%.0f QPS), device %.1f us, recall@1 %s" % (nq, host_s * 1e6, nq / host_s, dev_s * 1e6, hit))
    idx.close()
    return {"workload": name, "nb": w["nb"], "nc": w["nc"], "code_size": w["M"], "nprobe": w["nprobe"], "k": k, "rows": rows}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workloads", default="c5s,c5b")
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "r02_bench_c5_sweep.json"))
    args = ap.parse_args()
    import torch
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    batches = [1, 10, 100, 1000, 10000, 100000]
    reps_of = lambda nq: max(3, min(200, 200000 // max(nq, 1000) * 4))
    out = {"note": "BASELINE configs[4] at 1/50 scale, one B200; host_call = ivfhnsw_b200_search with pinned host buffers",
           "results": [sweep(n, torch, dev, batches, reps_of) for n in args.workloads.split(",")]}
    with open(args.out, "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
