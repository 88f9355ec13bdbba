"""TEST/BENCH INFRASTRUCTURE -- the CPU reference arm of bench.py.

    ref_worker.py <workdir> build    build + save the HNSW over the centroid file (reference code), nothing else
    ref_worker.py <workdir> run      load everything ONCE, fork cfg["nproc"] workers, each searches its slice of the
                                     queries sequentially; writes <workdir>/result.json (+ parity.npz)

Runs the COMPILED REFERENCE (oracle/_ref/libivfhnsw_ref.so): IndexIVF_HNSW[_Grouping]::build_quantizer, read,
rotate_quantizer and the sequential search loop of tests/test_ivfhnsw_sift1b.cpp:192-213.  No GPU code here.

The reference's search() is not re-entrant (member scratch) and it ships no threaded query loop, so the all-cores
figure comes from processes.  They are FORKED after read(): the 25 GB of a 1B-vector index are loaded once and shared
copy-on-write (nproc separate loads would not fit the host), the reference code itself stays unmodified.
Each worker also leaves, for the parity check of the GPU arm, what the reference answered for its slice together with
the coarse candidate lists its HNSW produced (the TRACE_CENTROIDS vectors, IndexIVF_HNSW.h:74-75).
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
    t0 = time.time()
    # loads info/edges when they exist, else constructs and saves them (IndexIVF_HNSW.cpp:40-65)
    r.L.ref_build_quantizer(r.h, p("centroids.fvecs"), p("hnsw.info"), p("hnsw.edges"), 16, int(w.get("efc", 120)))
    t_hnsw = time.time() - t0
    if mode == "build":
        return
    t0 = time.time()
    r.L.ref_load_pq(r.h, p("pq.pq"), p("norm.pq"), p("opq.mat") if w["opq"] else b"")
    r.set_do_opq(w["opq"])
    r.read(os.path.join(workdir, "x.index"))
    if w["opq"]:
        r.rotate_quantizer()
    r.set_params(w["nprobe"], w["max_codes"], w["efs"], w["pruning"])
    t_read = time.time() - t0
    Q = np.load(os.path.join(workdir, "queries.npy"))
    gt = np.load(os.path.join(workdir, "gt0.npy"))
    n, nproc, k = Q.shape[0], int(w["nproc"]), int(w["k"])
    steps, warmup = max(1, int(w.get("steps", 1))), int(w.get("warmup", 0))
    kr = int(w.get("k_recall", 100))
    start_at = time.time() + 1.0          # all workers enter their timed loop together
    pids = []
    for rank in range(nproc):
        pid = os.fork()
        if pid:
            pids.append(pid)
            continue
        try:
            lo, hi = rank * n // nproc, (rank + 1) * n // nproc
            Qs = np.ascontiguousarray(Q[lo:hi])
            if hi > lo:
                for _ in range(warmup):
                    r.search(Qs[: max(1, (hi - lo) // 4)], k, want_coarse=False)
                while time.time() < start_at:
                    pass
                t1 = time.perf_counter()
                for _ in range(steps):
                    r.search(Qs, k, want_coarse=False)
                sec = time.perf_counter() - t1
                # untimed extras: the answers + coarse lists for the parity check, recall@1/10/100
                D, L, cd, ci = r.search(Qs, k, want_coarse=True)
                Dr, Lr = r.search(Qs, kr, want_coarse=False)
                o = np.argsort(Dr, axis=1, kind="stable")
                Lr = np.take_along_axis(Lr, o, 1)
                hit = Lr == gt[lo:hi, None]
                rec = [int(hit[:, :m].any(1).sum()) for m in (1, 10, 100)]
                np.savez(os.path.join(workdir, "out_%d.npz" % rank), D=D, L=L, cd=cd, ci=ci)
            else:
                sec, rec = 0.0, [0, 0, 0]
            with open(os.path.join(workdir, "out_%d.json" % rank), "w") as f:
                json.dump({"n": (hi - lo) * steps, "nq": hi - lo, "sec": sec, "rec": rec}, f)
            os._exit(0)
        except BaseException as ex:       # never fall back into the parent's code path
            sys.stderr.write("ref worker %d failed: %r\n" % (rank, ex))
            os._exit(1)
    bad = 0
    for pid in pids:
        _, status = os.waitpid(pid, 0)
        if status != 0:
            bad += 1
            sys.stderr.write("ref worker pid %d: exit status %d signal %d\n" % (pid, status >> 8, status & 0x7f))
    os.dup2(saved, 1)
    if bad:
        raise SystemExit("ref_worker: %d workers failed" % bad)
    outs = [json.load(open(os.path.join(workdir, "out_%d.json" % rank))) for rank in range(nproc)]
    live = [o for o in outs if o["nq"]]
    parts = [np.load(os.path.join(workdir, "out_%d.npz" % rank)) for rank in range(nproc) if outs[rank]["nq"]]
    np.savez(os.path.join(workdir, "parity.npz"), **{key: np.concatenate([q[key] for q in parts]) for key in ("D", "L", "cd", "ci")})
    nq = sum(o["nq"] for o in live)
    res = dict(qps=sum(o["n"] / o["sec"] for o in live),            # every worker times its own sequential loop
               us_per_query=float(np.mean([o["sec"] / o["n"] * 1e6 for o in live])),
               ms_per_step=1e3 * max(o["sec"] for o in live) / steps,
               recall={"r1": sum(o["rec"][0] for o in live) / nq, "r10": sum(o["rec"][1] for o in live) / nq,
                       "r100": sum(o["rec"][2] for o in live) / nq},
               hnsw_s=t_hnsw, read_s=t_read, nq=nq, nproc=nproc)
    with open(os.path.join(workdir, "result.json"), "w") as f:
        json.dump(res, f)


if __name__ == "__main__":
    main()
