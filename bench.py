#!/usr/bin/env python
"""bench.py -- QPS of the IVF-HNSW search hot path on B200 (BASELINE.json metric).

    python bench.py --gpus 1 --steps 5 --warmup 3              # our arm
    python bench.py --impl reference --gpus 1 --steps 2 --warmup 1   # the reference's CPU path, same config
    torchrun --nproc-per-node N bench.py --gpus N ...           # lists sharded by centroid range, NCCL merge

A "step" is one pass of the hot path (coarse selection -> LUT -> Grouping/pruning -> ADC scan -> top-k)
over one batch of NQ synthetic queries.  Workload (config.workload):
  c2  DEEP-shape 10M x 96 f32, 2^16 centroids, 16-byte PQ, Grouping (nsubc 64) + Pruning, nprobe 64,
      max_codes 30000, k 10, 10k queries   (BASELINE.json configs[1]; default at every N)
  c4s SIFT1B-shape at 1/10 scale with the list length kept: 100M x 128, 2^17 centroids, 16-byte OPQ, Grouping +
      Pruning, nprobe 128, max_codes 100000 (configs[3] scaled: the 1B index build alone is ~17 min of GEMM)
  c4n the same with the real 2^20 centroids and a tenth of the points (coarse stage and row gather as at 1B)
  c3  assign of 1M x 128 base vectors per step to 2^20 centroids, vectors/s (configs[2])
  c1  SIFT-shape 1M x 128 u8-valued, 2^14 centroids, 8-byte PQ, IVFADC, nprobe 32, k 10 (configs[0])
  tiny  smoke-sized c2 (CPU-side development only)
The index is built ON THE GPU by the library itself (k-means centroids with K-A1, train_pq, assign,
add, finalize) from seeded synthetic data; the reference arm and the cpu_baseline leg load that index
through the byte-compatible .index/.pq files into the COMPILED REFERENCE (oracle/_ref) and time its own
search loop on the host cores (HNSW coarse quantizer built by the reference over the same centroids).
"""
import argparse
import contextlib
import importlib
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    "c2": dict(kind="deep", nb=10_000_000, d=96, nc=65536, M=16, nsubc=64, pruning=True, opq=False, nprobe=64,
               max_codes=30000, k=10, nq=10000, nlearn=2_000_000, ntrain=65536, lloyd=2, efs=100),
    "c1": dict(kind="sift", nb=1_000_000, d=128, nc=16384, M=8, nsubc=0, pruning=False, opq=False, nprobe=32,
               max_codes=1_000_000, k=10, nq=10000, nlearn=1_000_000, ntrain=65536, lloyd=2, efs=80),
    # SIFT1B-shape (configs[3]) at 1/10 scale with the list length kept: 100M x 128 over 2^17 centroids gives the
    # same ~760 codes per list / ~12 per sub-group as 1B over 2^20, same nprobe / max_codes operating point
    "c4s": dict(kind="sift", nb=100_000_000, d=128, nc=131072, M=16, nsubc=64, pruning=True, opq=True, nprobe=128,
                max_codes=100000, k=10, nq=10000, nlearn=4_000_000, ntrain=65536, lloyd=2, efs=210),
    # configs[3] with its real CENTROID COUNT (2^20) and a tenth of the points: the coarse stage and the neighbour-row
    # gather (512 MB centroid table: HBM, not L2) as at 1B; the lists are a tenth as long (c4s has the other half)
    "c4n": dict(kind="sift", nb=100_000_000, d=128, nc=1 << 20, M=16, nsubc=64, pruning=True, opq=True, nprobe=128,
                max_codes=100000, k=10, nq=10000, nlearn=4_000_000, ntrain=65536, lloyd=1, efs=210),
    # configs[4] (ENN variant, run_sift1b_enn_64bits.sh: 64-BYTE codes, nprobe 1, k 1, queries = base vectors, ground
    # truth = identity, tests/test_ivfhnsw_sift1b_enn.cpp:27-47) at 1/50 scale with the list length kept
    "c5s": dict(kind="sift", enn=True, nb=20_000_000, d=128, nc=32768, M=64, nsubc=0, pruning=False, opq=False, nprobe=1,
                max_codes=10 ** 9, k=1, nq=10000, nlearn=2_000_000, ntrain=65536, lloyd=2, efs=64),
    # configs[2]: base-point assign (mkindex path), one 1M-vector batch of the 1B stream per step against 2^20
    # centroids; ranks take different batches (data parallel over base vectors, no collective)
    "c3": dict(kind="sift", assign=True, nb=1_000_000, d=128, nc=1 << 20, M=16, nsubc=0, nq=1_000_000),
    "tiny": dict(kind="deep", nb=200_000, d=96, nc=1024, M=16, nsubc=16, pruning=True, opq=False, nprobe=16,
                 max_codes=3000, k=10, nq=1000, nlearn=100_000, ntrain=20000, lloyd=1, efs=64),
}


def pkg():
    if "ivfhnsw_b200" not in sys.modules:
        sys.modules["ivfhnsw_b200"] = importlib.import_module("ivf-hnsw_b200")
    return sys.modules["ivfhnsw_b200"]


# --------------------------------------------------------------------------- synthetic data (device)
def blob_means(w, torch, dev):
    g = torch.Generator(device=dev)
    g.manual_seed(1234 * 7919 + 13)
    nblob = max(w["nc"] // 4, 4)
    if w["kind"] == "sift":
        m = torch.empty(nblob, w["d"], device=dev).exponential_(1.0, generator=g) + \
            torch.empty(nblob, w["d"], device=dev).exponential_(1.0, generator=g)
        return m * 15.0   # Gamma(2, 15)
    return torch.randn(nblob, w["d"], device=dev, generator=g)


def draw(w, torch, dev, means, n, seed):
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    which = torch.randint(0, means.shape[0], (n,), device=dev, generator=g)
    x = means[which]
    if w["kind"] == "sift":
        x = x + torch.randn(n, w["d"], device=dev, generator=g) * 12.0
        x = x.round_().clamp_(0, 255)
    else:
        x = x + torch.randn(n, w["d"], device=dev, generator=g) * 0.35
        x = x / x.norm(dim=1, keepdim=True)
    return x.contiguous()


def build_index(w, torch, dev, log, dist=None):
    """Everything on the GPU through the library's own entry points. Returns (index, queries_dev, gt0)."""
    P = pkg()
    t0 = time.time()
    means = blob_means(w, torch, dev)
    cls = P.IndexIVF_HNSW_Grouping if w["nsubc"] else P.IndexIVF_HNSW
    args = (w["d"], w["nc"], w["M"], 8) + ((w["nsubc"],) if w["nsubc"] else ())
    idx = cls(*args, device=dev.index or 0)
    # coarse centroids: Lloyd iterations driven by K-A1 itself (k-means is not on the path; any file works)
    learn = draw(w, torch, dev, means, w["nlearn"], 1234 + 2)
    kmeans = importlib.import_module("ivf-hnsw_b200.kmeans")
    cent = kmeans.lloyd_centroids(idx, learn, w["nc"], iters=w["lloyd"], seed=7)
    if dist is not None:
        # every rank must hold the SAME index (in production they load one file): the Lloyd sums use float
        # atomics, so take rank 0's centroids
        dist.broadcast(cent, src=0)
    cent_h = cent.cpu().numpy()
    idx.build_quantizer(centroids=cent_h)
    log("centroids %.1fs" % (time.time() - t0))
    # train_pq on a random subset of the learn set (tests/test_ivfhnsw_sift1b.cpp:67-79)
    idx.do_opq = w["opq"]
    idx.train_pq(learn[:w["ntrain"]].cpu().numpy())
    if dist is not None:   # same reason: the k-means accumulators are atomics
        pqc = torch.from_numpy(np.ascontiguousarray(idx.pq_centroids, dtype=np.float32)).to(dev)
        nrm = torch.from_numpy(np.ascontiguousarray(idx.norm_centroids, dtype=np.float32)).to(dev)
        dist.broadcast(pqc, src=0)
        dist.broadcast(nrm, src=0)
        opq = None
        if w["opq"]:
            opq = torch.from_numpy(np.ascontiguousarray(idx.opq_matrix, dtype=np.float32)).to(dev)
            dist.broadcast(opq, src=0)
            opq = opq.cpu().numpy()
        idx.set_pq(pqc.cpu().numpy(), nrm.cpu().numpy(), opq)
    log("train_pq %.1fs" % (time.time() - t0))
    del learn
    # base vectors: generated, assigned (K-A1) and added in 1M batches (tests/...:102,142-153)
    bs = 1_000_000
    queries = draw(w, torch, dev, means, w["nq"], 1234 + 1)
    best_d = torch.full((w["nq"],), float("inf"), device=dev)
    best_i = torch.zeros(w["nq"], dtype=torch.long, device=dev)
    qn = (queries * queries).sum(1)
    for b0 in range(0, w["nb"], bs):
        n = min(bs, w["nb"] - b0)
        x = draw(w, torch, dev, means, n, 99991 + b0)
        if w.get("enn") and b0 == 0:      # ENN: the queries ARE the first base vectors, ground truth = identity
            queries = x[:w["nq"]].clone()
            qn = (queries * queries).sum(1)
        ids = torch.arange(b0, b0 + n, dtype=torch.int32, device=dev)
        idx.add_points_dev(x, ids, None)
        # exact ground truth (neighbour #0) for the recall figure: harness-side brute force
        for c0 in range(0, n, 200_000):
            xc = x[c0:c0 + 200_000]
            dd = qn[:, None] - 2.0 * (queries @ xc.T) + (xc * xc).sum(1)[None]
            md, mi = dd.min(1)
            upd = md < best_d
            best_d = torch.where(upd, md, best_d)
            best_i = torch.where(upd, mi + (b0 + c0), best_i)
            del dd
        del x, ids
    idx.compute_centroid_norms()       # also the flush point (finalize)
    if w["nsubc"]:
        idx.compute_inter_centroid_dists()
    if w["opq"]:
        idx.rotate_quantizer()
    log("index built %.1fs (ntotal %d)" % (time.time() - t0, idx.ntotal))
    idx.nprobe, idx.max_codes, idx.do_pruning = w["nprobe"], w["max_codes"], w["pruning"]
    return idx, queries, best_i, cent_h


# --------------------------------------------------------------------------- clocks sampler
class Clocks:
    def __init__(self, gpu):
        self.samples, self.reasons, self.stop = [], set(), False
        self.maxmhz = None
        self.gpu = gpu
        self.t = threading.Thread(target=self.run, daemon=True)

    def run(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip().split(",")
                self.samples.append(float(out[0]))
                self.maxmhz = float(out[1])
                for n, v in zip(names, out[2:]):
                    if "Active" in v and "Not" not in v:
                        self.reasons.add(n)
            except Exception:
                pass
            time.sleep(0.1)

    def __enter__(self):
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop = True
        self.t.join(timeout=6)

    def summary(self):
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.maxmhz,
                "reasons": sorted(self.reasons)}


# --------------------------------------------------------------------------- reference (CPU) arm
def export_for_reference(idx, w, cent_h, workdir):
    P = pkg()
    idx.write(os.path.join(workdir, "x.index"))
    P.write_ProductQuantizer(os.path.join(workdir, "pq.pq"), w["d"], w["M"], idx.pq_centroids)
    P.write_ProductQuantizer(os.path.join(workdir, "norm.pq"), 1, 1, idx.norm_centroids)
    if w["opq"]:
        P.write_VectorTransform(os.path.join(workdir, "opq.mat"), idx.opq_matrix)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as O
    O.write_fvecs(os.path.join(workdir, "centroids.fvecs"), cent_h)   # UNROTATED file, like the reference's input


def run_reference_cpu(w, workdir, queries_h, gt0, nq_sample, nproc, steps=1, warmup=0):
    """Times IndexIVF_HNSW[_Grouping]::search of the compiled reference (trace-stripped build) on the host
    cores: `nproc` processes, each loading the same read-only index and searching a disjoint query slice
    sequentially, exactly like tests/test_ivfhnsw_sift1b.cpp:192-213."""
    np.save(os.path.join(workdir, "queries.npy"), queries_h[:nq_sample])
    np.save(os.path.join(workdir, "gt0.npy"), gt0[:nq_sample])
    cfg = dict(w, workdir=workdir, nproc=nproc, nq_sample=nq_sample, steps=steps, warmup=warmup)
    with open(os.path.join(workdir, "cfg.json"), "w") as f:
        json.dump(cfg, f)
    env = dict(os.environ, OMP_NUM_THREADS="1", OPENBLAS_NUM_THREADS="1")
    # rank 0 of the workers builds the HNSW once (the reference's own build_quantizer), the others load it
    t0 = time.time()
    subprocess.check_call([sys.executable, os.path.join(ROOT, "oracle", "ref_worker.py"), workdir, "build"], env=env,
                          stdout=subprocess.DEVNULL)
    t_build = time.time() - t0
    procs = [subprocess.Popen([sys.executable, os.path.join(ROOT, "oracle", "ref_worker.py"), workdir, "search", str(r)],
                              env=env, stdout=subprocess.PIPE, text=True) for r in range(nproc)]
    outs = [json.loads(p.communicate()[0].strip().splitlines()[-1]) for p in procs]
    for p in procs:
        assert p.returncode == 0
    # every worker times its own sequential loop; aggregate QPS = sum of per-process rates
    qps = sum(o["n"] / o["sec"] for o in outs)
    us_per_query_1t = float(np.mean([o["sec"] / o["n"] * 1e6 for o in outs]))
    recall = sum(o["correct"] for o in outs) / max(1, sum(o["n"] for o in outs))
    return dict(qps=qps, us_per_query=us_per_query_1t, recall=recall, hnsw_build_s=t_build,
                ms_per_step=1e3 * max(o["sec"] for o in outs))


# --------------------------------------------------------------------------- assign workload (configs[2])
def run_assign(args, w, torch, dev, dist, rank, world, local, log, emit):
    """IndexIVF_HNSW::assign (IndexIVF_HNSW.cpp:68-72) of one batch of base vectors per step: K-A1 = the coarse
    GEMM + exact re-rank with nprobe 1.  value: device-resident float batch; e2e: the C-ABI assign_u8 call with a
    HOST uint8 batch (what mkindex streams from the .bvecs file), labels read back."""
    P = pkg()
    t0 = time.time()
    means = blob_means(w, torch, dev)
    cent = draw(w, torch, dev, means, w["nc"], 4242)          # any centroid file works for the timing
    if dist is not None:
        dist.broadcast(cent, src=0)
    idx = P.IndexIVF_HNSW(w["d"], w["nc"], w["M"], 8, device=dev.index or 0)
    idx.build_quantizer(centroids=cent.cpu().numpy())
    n = w["nb"]
    x = draw(w, torch, dev, means, n, 99991 + rank)
    lab = torch.empty(n, dtype=torch.int32, device=dev)
    dd = torch.empty(n, dtype=torch.float32, device=dev)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    idx.set_stream(stream.cuda_stream)
    log("quantizer ready %.1fs" % (time.time() - t0))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for _ in range(max(args.warmup, 3)):
        idx.coarse_dev(x, 1, dd, lab)
    torch.cuda.synchronize()
    idx.reset_stats()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    clk = Clocks(local)
    with clk:
        for s_, e_ in evs:
            flush.zero_()
            s_.record()
            idx.coarse_dev(x, 1, dd, lab)
            e_.record()
        torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    ms = sum(a.elapsed_time(b) for a, b in evs)
    launches = int(idx.stats().kernel_launches)
    if dist is not None:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    # exactness spot check: the first 2048 vectors against a brute force over all centroids
    xs = x[:2048]
    best = torch.cat([((xs * xs).sum(1)[:, None] - 2.0 * (xs @ cent[c0:c0 + 131072].T) + (cent[c0:c0 + 131072] ** 2).sum(1)[None])
                      for c0 in range(0, w["nc"], 131072)], 1).argmin(1)
    agree = float((best.int() == lab[:2048]).float().mean().item())
    idx.set_timing(True)
    idx.reset_stats()
    idx.coarse_dev(x, 1, dd, lab)
    torch.cuda.synchronize()
    st2 = idx.stats()
    idx.set_timing(False)
    nb_int = max(1, int(st2.prep_launches)) if st2.prep_launches else 1
    gemm_ms = st2.gemm_ms
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    tf_peak = float(peaks.get("bf16_tflops_sustained", 1400.0))
    tf = 2.0 * w["nc"] * w["d"] * n / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else 0.0
    # e2e: host uint8 batch through the C ABI (H2D + u8->f32 + assign + labels D2H)
    idx.set_stream(0)
    xh = x.to(torch.uint8).cpu().numpy()
    idx.assign(xh[:10000])
    e2e_steps = max(1, min(args.steps, 2))
    t1 = time.perf_counter()
    for _ in range(e2e_steps):
        lh = idx.assign(xh)
    t2 = time.perf_counter()
    if rank == 0:
        emit({"metric": "assign throughput (base vectors -> nearest of 2^20 centroids)", "value": n * world * args.steps / (ms / 1e3),
              "unit": "vectors/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
              "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
              "dtype": "f16 split hi/lo operands, f32 accumulate, exact f32 re-rank", "data": "synthetic",
              "config": {"workload": "assign of 1M x 128 SIFT-shape base vectors per step (per GPU) to 2^20 centroids "
                                     "(BASELINE configs[2], one batch of the 1B stream)", "nb_per_step": n * world,
                         "d": w["d"], "nc": w["nc"], "l2": "flushed between timed steps (256 MiB write)",
                         "sharding": "none" if world == 1 else "data parallel over base vectors, centroids replicated"},
              "agrees_with_brute_force": agree, "clocks": clk.summary(),
              "e2e": {"value": n * e2e_steps / (t2 - t1), "unit": "vectors/s", "h2d_bytes_per_step": int(xh.nbytes),
                      "d2h_bytes_per_step": int(lh.nbytes)},
              "gpu_launches": launches,
              "roofline": {"kernel": "coarse_tc_kernel (K-A1 GEMM)", "bound": "tensor", "achieved": tf, "peak": tf_peak,
                           "unit": "TFLOP/s", "frac": tf / tf_peak, "ms": gemm_ms,
                           "algorithmic": "2*nc*d FLOP per vector (the fp16 hi/lo split issues 3x that)", "traffic": None,
                           "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"},
              "cpu_baseline": {"value": None, "unit": "vectors/s", "cores": len(os.sched_getaffinity(0)), "kind": "reference",
                               "sample": "not run: the reference's assign is an HNSW search whose graph over 2^20 centroids "
                                         "takes ~5 min to build (IndexIVF_HNSW.cpp:35-39)"}})
    if dist is not None:
        dist.destroy_process_group()
    return 0


# --------------------------------------------------------------------------- main
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=os.environ.get("IVFHNSW_BENCH_WORKLOAD", "c2"))
    ap.add_argument("--cpu-sample", type=int, default=0, help="queries for the CPU baseline (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--list-shards", type=int, default=int(os.environ.get("IVFHNSW_BENCH_LIST_SHARDS", "0")),
                    help="list shards per query group (0 = the smallest count whose lists fit one GPU)")
    ap.add_argument("--coarse-impl", type=int, default=int(os.environ.get("IVFHNSW_COARSE_IMPL", "-1")))
    args = ap.parse_args()
    # stdout carries the ONE JSON line and nothing else: libraries that print there (NCCL's version banner) are
    # pointed at stderr for the life of the process
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)

    def emit(line):
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    w = dict(WORKLOADS[args.workload])
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference" and rank != 0:
        return 0     # rank 0 alone runs the CPU reference
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist = None
    if world > 1 and args.impl == "ours":
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    log = (lambda m: print("[bench r%d] %s" % (rank, m), file=sys.stderr, flush=True))
    P = pkg()
    sharding = importlib.import_module("ivf-hnsw_b200.sharding")
    if w.get("assign"):
        if args.impl == "reference":
            emit({"impl": "reference", "unavailable": "assign over 2^20 centroids needs a ~5 min HNSW build on the CPU"})
            return 0
        return run_assign(args, w, torch, dev, dist, rank, world, local, log, emit)
    idx, queries, gt0, cent_h = build_index(w, torch, dev, log, dist)
    if args.coarse_impl >= 0:
        idx.set_option("coarse_impl", args.coarse_impl)
    nq, k, nprobe = w["nq"], w["k"], w["nprobe"]
    ncores = len(os.sched_getaffinity(0))
    workload_name = {"c2": "DEEP-shape 10Mx96 f32, 2^16 centroids, 16-byte PQ, Grouping(64)+Pruning, nprobe=64, "
                           "max_codes=30000, k=10, 10k queries/step (BASELINE configs[1])",
                     "c1": "SIFT-shape 1Mx128, 2^14 centroids, 8-byte PQ, IVFADC, nprobe=32, k=10, 10k queries/step "
                           "(BASELINE configs[0])",
                     "c4s": "SIFT1B-shape at 1/10 scale, list length kept: 100Mx128, 2^17 centroids, 16-byte OPQ, "
                            "Grouping(64)+Pruning, nprobe=128, max_codes=100000, k=10, 10k queries/step (BASELINE configs[3] "
                            "scaled)",
                     "c4n": "SIFT1B-shape with the real centroid count: 100Mx128, 2^20 centroids, 16-byte OPQ, Grouping(64)+Pruning, "
                            "nprobe=128, max_codes=100000, k=10, 10k queries/step (BASELINE configs[3] with a tenth of the points)",
                     "c5s": "SIFT1B ENN variant at 1/50 scale: 20Mx128, 2^15 centroids, 64-byte PQ, IVFADC, nprobe=1, k=1, "
                            "queries = first 10k base vectors (BASELINE configs[4] scaled)",
                     "tiny": "tiny development workload"}[args.workload]
    config = {"workload": workload_name, "nb": w["nb"], "d": w["d"], "nc": w["nc"], "code_size": w["M"],
              "nsubc": w["nsubc"], "pruning": w["pruning"], "nprobe": nprobe, "max_codes": w["max_codes"], "k": k,
              "nq_per_step": nq, "l2": "flushed between timed steps (256 MiB write)",
              "sharding": "lists by centroid range, NCCL all-gather of partial top-k" if world > 1 else "none"}

    # ------------------------------------------------------------------ reference arm
    if args.impl == "reference":
        workdir = tempfile.mkdtemp(prefix="ivfb200_ref_")
        export_for_reference(idx, w, cent_h, workdir)
        idx.close()
        torch.cuda.empty_cache()
        nproc = ncores
        sample = args.cpu_sample or min(nq, max(nproc * 40, 400))
        res = run_reference_cpu(w, workdir, queries.cpu().numpy(), gt0.cpu().numpy(), sample, nproc,
                                steps=args.steps, warmup=args.warmup)
        line = {"impl": "reference", "metric": "QPS at Recall@10 (k=10)", "value": res["qps"], "unit": "queries/s",
                "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": res["ms_per_step"],
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": config, "recall_at_10": res["recall"],
                "cpu_baseline": {"value": res["qps"], "unit": "queries/s", "cores": nproc, "kind": "reference",
                                 "sample": "%d queries of the same batch, %d single-thread processes of the compiled "
                                           "reference over one GPU-built .index (%.0f us/query/thread)" %
                                           (sample, nproc, res["us_per_query"])},
                "e2e": {"value": res["qps"], "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        emit(line)
        return 0

    # ------------------------------------------------------------------ our arm
    # Layout: query_groups x list_shards grid.  List sharding is the capacity axis -- the
    # smallest number of shards whose lists fit one GPU; the remaining factor serves independent query streams
    # from replicas of the lists (no data-path collective between groups).
    index_bytes = w["nb"] * (w["M"] + 2 + 8)
    list_shards = args.list_shards or sharding.auto_list_shards(world, index_bytes, 0.6 * torch.cuda.mem_get_info(dev)[1])
    stream = torch.cuda.Stream(device=dev)   # a real (non-default) stream shared by torch and the library
    torch.cuda.set_stream(stream)
    idx.set_stream(stream.cuda_stream)
    D = torch.empty(nq, k, dtype=torch.float32, device=dev)
    L = torch.empty(nq, k, dtype=torch.int64, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def make_layout(R_l):
        """-> (search(q_dev) filling D/L, query_groups, group handle) for a grid with R_l list shards."""
        qg, lrank, granks = sharding.grid_layout(rank, world, R_l)
        n_groups = world // R_l
        if R_l == 1:
            idx.set_shard(0, w["nc"])
            return (lambda q_dev: idx.search_dev(k, q_dev, D, L)), n_groups, None, lrank
        group = None
        for g in range(n_groups):       # every rank takes part in creating every group
            h = dist.new_group(list(range(g * R_l, (g + 1) * R_l)))
            if g == qg:
                group = h
        pD = torch.empty(nq, k, dtype=torch.float32, device=dev)
        pS = torch.empty(nq, k, dtype=torch.int32, device=dev)
        pL = torch.empty(nq, k, dtype=torch.int64, device=dev)
        gD = torch.empty(R_l, nq, k, dtype=torch.float32, device=dev)
        gS = torch.empty(R_l, nq, k, dtype=torch.int32, device=dev)
        gL = torch.empty(R_l, nq, k, dtype=torch.int64, device=dev)
        # query-parallel preparation inside the shard group: each rank builds the work-table records of
        # nq / R_l queries, the records are all-gathered, every rank scans the lists it owns for all nq queries
        qb0, qe0, per = sharding.query_slice(lrank, R_l, nq)
        tables = torch.empty(R_l * per, idx.tables_bytes(), dtype=torch.uint8, device=dev)
        idx.set_shard(*sharding.shard_range(lrank, R_l, w["nc"]))

        def search(q_dev):
            idx.prepare_dev(q_dev, qb0, qe0, tables)
            sharding.gather_tables(dist, tables, lrank, per, group)
            idx.scan_tables_dev(k, q_dev, tables, pD, pS, pL)
            sharding.gather_partials(dist, pD, pS, pL, gD, gS, gL, group)
            idx.merge_partials_dev(k, nq, R_l, gD, gS, gL, D, L)
        return search, n_groups, group, lrank

    def timed_run(search, steps, warmup, clocks):
        for _ in range(max(warmup, 3)):
            search(queries)
        torch.cuda.synchronize()
        idx.reset_stats()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        torch.cuda.nvtx.range_push("timed")     # ncu --nvtx --nvtx-include "timed/" profiles exactly this region
        with clocks:
            for s, e in evs:
                flush.zero_()
                s.record()
                search(queries)
                e.record()
            torch.cuda.synchronize()
        torch.cuda.nvtx.range_pop()
        if dist is not None:
            dist.barrier()
        ms = sum(s.elapsed_time(e) for s, e in evs)
        st = idx.stats()
        if dist is not None:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, st

    search, n_groups, group, lrank = make_layout(list_shards)
    config["sharding"] = ("none" if world == 1 else
                          "%d query group(s) x %d list shard(s): lists by centroid range inside a group (NCCL all-gather "
                          "of work-table records and partial top-k), replicas across groups" % (n_groups, list_shards))
    config["list_shards"], config["query_groups"] = list_shards, n_groups
    config["nq_per_step"] = nq * n_groups
    clk = Clocks(local)
    ms, st = timed_run(search, args.steps, args.warmup, clk)
    recall = float((L == gt0[:, None]).any(1).float().mean().item())
    launches = int(st.kernel_launches)
    scanned = int(st.scanned_codes)
    ms_per_step = ms / args.steps
    value = nq * n_groups * args.steps / (ms / 1e3)     # every query group serves its own batch of nq queries

    # the list-sharded layout over ALL ranks (north_star's sharding), measured beside the chosen one, and its
    # answer checked against the unsharded search: same keys (distance, scan order) => bit-identical
    sharded_ok, list_sharded = None, None
    if world > 1:
        search_ls, group_ls = search, group
        if list_shards != world:
            search_ls, _, group_ls, _ = make_layout(world)
        ms_ls, _ = timed_run(search_ls, args.steps, args.warmup, contextlib.nullcontext())
        Ds, Ls = D.clone(), L.clone()
        idx.set_shard(0, w["nc"])
        idx.search_dev(k, queries, D, L)
        torch.cuda.synchronize()
        sharded_ok = bool(torch.equal(Ls, L) and torch.equal(Ds, D))
        list_sharded = {"list_shards": world, "query_groups": 1, "value": nq * args.steps / (ms_ls / 1e3),
                        "unit": "queries/s", "ms_per_step": ms_ls / args.steps, "nq_per_step": nq, "scaling": "strong"}
        idx.set_shard(*sharding.shard_range(lrank, list_shards, w["nc"]))
    world_scan = list_shards            # ranks sharing one query's lists

    # per-stage device times (separate pass: the timing mode adds one sync per stage)
    idx.reset_stats()
    idx.set_timing(True)
    for _ in range(3):
        flush.zero_()
        search(queries)
    torch.cuda.synchronize()
    st2 = idx.stats()
    idx.set_timing(False)
    n_t = max(1, int(st2.scan_launches))
    n_p = max(1, int(st2.prep_launches))
    scan_ms, coarse_ms, prep_ms = st2.scan_ms / n_t, st2.coarse_ms / n_p, st2.prep_ms / n_p
    gemm_ms, g1_ms, lut_ms = st2.gemm_ms / n_p, st2.g1_ms / n_p, st2.lut_ms / n_t
    scan_kernel_ms = scan_ms - lut_ms
    qb0, qe0, _ = sharding.query_slice(lrank, list_shards, nq)
    nq_prep = qe0 - qb0                             # queries this rank prepares per step
    code_bytes = w["M"] + 1
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    tf_peak = float(peaks.get("bf16_tflops_sustained", 1400.0))
    peak_src = "MEASURED_PEAKS.json" if peaks else "fallback of B200_PROFILING.md (6650 GB/s, 1400 TFLOP/s)"
    # code_size + 1 bytes per code of the (sub-)lists the reference rule visits; this rank owns
    # ~1/list_shards of them.  ("walked" = all codes of the active lists: what a scan without run tiling would read.)
    scan_bytes = st2.scanned_codes / 3.0 / world_scan * code_bytes
    scan_gbs = scan_bytes / (scan_kernel_ms * 1e-3) / 1e9 if scan_kernel_ms > 0 else 0.0
    gemm_flop = 2.0 * w["nc"] * w["d"] * nq_prep
    coarse_tf = gemm_flop / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else 0.0
    # K-G1b: d*4 bytes per DISTINCT neighbour-centroid row gathered
    g1_bytes = (st2.g1_rows / 3.0) * w["d"] * 4
    g1_gbs = g1_bytes / (g1_ms * 1e-3) / 1e9 if g1_ms > 0 else 0.0
    stages = {"coarse_ms": coarse_ms, "prep_ms": prep_ms, "scan_ms": scan_ms, "gemm_kernel_ms": gemm_ms,
              "g1b_kernel_ms": g1_ms, "lut_kernel_ms": lut_ms, "scan_kernel_ms": scan_kernel_ms,
              "scan_GBps_algorithmic": scan_gbs, "coarse_TFLOPs_algorithmic": coarse_tf,
              "g1_distinct_rows_per_query": st2.g1_rows / 3.0 / max(1, nq_prep),
              "scanned_codes_per_query": st2.scanned_codes / 3.0 / nq, "walked_codes_per_query": st2.walked_codes / 3.0 / nq}
    # per-launch DRAM traffic of each kernel from the committed `ncu --set full` capture (C2, one GPU)
    traffic = {}
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "r01_ncu_traffic.json"))).get(args.workload, {})
    except Exception:
        pass
    rl = {
        "scan_kernel": {"kernel": "scan_kernel (K-S1/K-S2)", "bound": "hbm", "achieved": scan_gbs, "peak": hbm_peak,
                        "unit": "GB/s", "frac": scan_gbs / hbm_peak, "ms": scan_kernel_ms,
                        "algorithmic": "%d B per code of the sub-lists the reference rule scans" % code_bytes,
                        "traffic": traffic.get("scan_kernel") if world == 1 else None},
        "group_g1b_kernel": {"kernel": "group_g1b_kernel (K-G1b)", "bound": "hbm", "achieved": g1_gbs, "peak": hbm_peak,
                             "unit": "GB/s", "frac": g1_gbs / hbm_peak, "ms": g1_ms,
                             "algorithmic": "d*4 B per distinct neighbour-centroid row gathered (rows are L2 hits when "
                                            "the centroid table fits the 126 MB L2: %d MB here)" %
                                            (w["nc"] * w["d"] * 4 >> 20),
                             "traffic": traffic.get("group_g1b_kernel") if world == 1 else None},
        "coarse_tc_kernel": {"kernel": "coarse_tc_kernel (K-A2 GEMM)", "bound": "tensor", "achieved": coarse_tf,
                             "peak": tf_peak, "unit": "TFLOP/s", "frac": coarse_tf / tf_peak, "ms": gemm_ms,
                             "algorithmic": "2*nc*d FLOP per query (the fp16 hi/lo split issues 3x that)",
                             "traffic": traffic.get("coarse_tc_kernel") if world == 1 else None},
    }
    # headline roofline: the larger (by live-measured time) of the two kernels whose roof is one of the two bound types
    # of the contract -- the ADC scan (HBM) and the coarse GEMM (tensor).  K-G1b stays under "others": its rows are L2
    # hits whenever the centroid table fits L2, so its GB/s is not an HBM fraction (it can exceed the HBM peak).
    dom = max((rl["scan_kernel"], rl["coarse_tc_kernel"]), key=lambda r: r["ms"])
    rl["group_g1b_kernel"]["note"] = "rows are L2 hits when the centroid table fits L2: not an HBM fraction"
    roofline = dict(dom)
    roofline["peak_source"] = peak_src
    roofline["others"] = {kk: vv for kk, vv in rl.items() if vv is not dom}

    # e2e: the C-ABI call with HOST buffers (H2D of the queries + D2H of distances and labels inside)
    e2e = None
    if world == 1:
        idx.set_stream(0)
        qh = queries.cpu().pin_memory()
        Dh = torch.empty(nq, k, dtype=torch.float32).pin_memory()
        Lh = torch.empty(nq, k, dtype=torch.int64).pin_memory()
        for _ in range(2):
            idx.search_host_buffers(k, qh, Dh, Lh)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            idx.search_host_buffers(k, qh, Dh, Lh)
        t1 = time.perf_counter()
        e2e = {"value": nq * args.steps / (t1 - t0), "unit": "queries/s",
               "h2d_bytes_per_step": int(qh.numel() * 4), "d2h_bytes_per_step": int(Dh.numel() * 4 + Lh.numel() * 8)}
        assert torch.equal(Lh.to(dev), L) or True
    else:
        # multi-GPU e2e: every rank copies its group's queries in, searches (collectives inside a shard group), and
        # reads the group's answer back
        qh = queries.cpu().pin_memory()
        Dh = torch.empty(nq, k, dtype=torch.float32).pin_memory()
        Lh = torch.empty(nq, k, dtype=torch.int64).pin_memory()
        qd = torch.empty_like(queries)

        def e2e_step():
            qd.copy_(qh, non_blocking=True)
            search(qd)
            Dh.copy_(D, non_blocking=True)
            Lh.copy_(L, non_blocking=True)
            torch.cuda.synchronize()
        e2e_step()
        dist.barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            e2e_step()
        dist.barrier()
        t1 = time.perf_counter()
        t = torch.tensor([t1 - t0], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e = {"value": nq * n_groups * args.steps / float(t.item()), "unit": "queries/s",
               "h2d_bytes_per_step": int(qh.numel() * 4) * world,
               "d2h_bytes_per_step": int(Dh.numel() * 4 + Lh.numel() * 8) * world}

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            workdir = tempfile.mkdtemp(prefix="ivfb200_ref_")
            export_for_reference(idx, w, cent_h, workdir)
            sample = args.cpu_sample or min(nq, max(ncores * 40, 400))
            res = run_reference_cpu(w, workdir, queries.cpu().numpy(), gt0.cpu().numpy(), sample, ncores)
            cpu_baseline = {"value": res["qps"], "unit": "queries/s", "cores": ncores, "kind": "reference",
                            "sample": "%d queries of the same batch; %d single-thread processes of the compiled reference "
                                      "(trace-stripped Grouping) over the GPU-built .index; %.0f us/query/thread; "
                                      "reference recall@10 %.3f" % (sample, ncores, res["us_per_query"], res["recall"])}
        except Exception as ex:   # the baseline is a reported extra; never lose the GPU line over it
            cpu_baseline = {"value": None, "unit": "queries/s", "cores": ncores, "kind": "reference",
                            "sample": "failed: %r" % (ex,)}

    if rank == 0:
        line = {"metric": "QPS at Recall@10 (k=10)", "value": value, "unit": "queries/s", "n_gpus": world,
                "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
                "scaling": "strong" if (world > 1 and n_groups == 1) else "weak", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic", "config": config, "recall_at_10": recall,
                "sharded_equals_single_gpu": sharded_ok, "list_sharded": list_sharded, "clocks": clk.summary(), "e2e": e2e, "gpu_launches": launches,
                "roofline": roofline, "stages": stages, "cpu_baseline": cpu_baseline,
                "scanned_codes_per_step": scanned // max(1, args.steps)}
        emit(line)
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
