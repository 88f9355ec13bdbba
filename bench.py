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

GEN_VERSION = 4   # bump when the synthetic generator changes (part of the index-cache key)

WORKLOADS = {
    "c2": dict(kind="deep", nb=10_000_000, d=96, nc=65536, M=16, nsubc=64, pruning=True, opq=False, nprobe=64,
               max_codes=30000, k=10, nq=10000, nlearn=2_000_000, ntrain=65536, lloyd=2, efs=100),
    "c1": dict(kind="sift", nb=1_000_000, d=128, nc=16384, M=8, nsubc=0, pruning=False, opq=False, nprobe=32,
               max_codes=1_000_000, k=10, nq=10000, nlearn=1_000_000, ntrain=65536, lloyd=2, efs=80),
    # BASELINE configs[3], the configuration the metric is quoted on: SIFT1B-shape 1B x 128, 2^20 centroids, 16-byte
    # OPQ, Grouping(64) + Pruning, nprobe 128, max_codes 100 000, k 10 (examples/run_sift1b_grouping_OPQ.sh:38-47)
    "c4": dict(kind="sift", nb=1_000_000_000, d=128, nc=1 << 20, M=16, nsubc=64, pruning=True, opq=True, nprobe=128,
               max_codes=100000, k=10, nq=10000, nlearn=16_000_000, ntrain=65536, lloyd=2, efs=210, gt_queries=2000),
    # the same at 1/10 scale with the list length kept: 100M x 128 over 2^17 centroids gives the same ~760 codes per
    # list / ~12 per sub-group as 1B over 2^20, same nprobe / max_codes operating point (development workload)
    "c4s": dict(kind="sift", nb=100_000_000, d=128, nc=131072, M=16, nsubc=64, pruning=True, opq=True, nprobe=128,
                max_codes=100000, k=10, nq=10000, nlearn=4_000_000, ntrain=65536, lloyd=2, efs=210),
    # configs[3] with its real CENTROID COUNT (2^20) and a tenth of the points (development workload)
    "c4n": dict(kind="sift", nb=100_000_000, d=128, nc=1 << 20, M=16, nsubc=64, pruning=True, opq=True, nprobe=128,
                max_codes=100000, k=10, nq=10000, nlearn=4_000_000, ntrain=65536, lloyd=1, efs=210),
    # configs[4] (ENN variant, run_sift1b_enn_64bits.sh: 64-BYTE codes, nprobe 1, k 1, queries = base vectors, ground
    # truth = identity, tests/test_ivfhnsw_sift1b_enn.cpp:27-47) at 1/50 scale with the list length kept
    "c5s": dict(kind="sift", enn=True, nb=20_000_000, d=128, nc=32768, M=64, nsubc=0, pruning=False, opq=False, nprobe=1,
                max_codes=10 ** 9, k=1, nq=10000, nlearn=2_000_000, ntrain=65536, lloyd=2, efs=64),
    # the "64 bits" reading of the same script name: 8-byte codes
    "c5b": dict(kind="sift", enn=True, nb=20_000_000, d=128, nc=32768, M=8, nsubc=0, pruning=False, opq=False, nprobe=1,
                max_codes=10 ** 9, k=1, nq=10000, nlearn=2_000_000, ntrain=65536, lloyd=2, efs=64),
    # configs[2]: base-point assign (mkindex path), one 1M-vector batch of the 1B stream per step against 2^20
    # centroids; ranks take different batches (data parallel over base vectors, no collective)
    "c3": dict(kind="sift", assign=True, nb=1_000_000, d=128, nc=1 << 20, M=16, nsubc=0, nq=1_000_000),
    "tiny": dict(kind="sift", nb=300_000, d=128, nc=1024, M=16, nsubc=16, pruning=True, opq=True, nprobe=16,
                 max_codes=3000, k=10, nq=1000, nlearn=100_000, ntrain=20000, lloyd=1, efs=64),
}
DEFAULT_WORKLOAD = "c4"


def pkg():
    if "ivfhnsw_b200" not in sys.modules:
        sys.modules["ivfhnsw_b200"] = importlib.import_module("ivf-hnsw_b200")
    return sys.modules["ivfhnsw_b200"]


# --------------------------------------------------------------------------- synthetic data (device)
class Gen:
    """Seeded synthetic base / query / learn vectors, generated on the device chunk by chunk (never stored: a chunk is
    a pure function of its seed, so the bounded-pass Grouping ingest can re-draw it like the reference re-reads the
    base file, tests/test_ivfhnsw_grouping_sift1b.cpp:143-171).

    SIFT-shape (d=128, byte-valued): nc/4 blobs whose means are drawn around nblob/8 super-means (neighbouring blobs
    exist, as in real descriptors); a point is its blob mean + a RANK-12 component (12 Gaussian factors through a
    global 12 x d matrix, dims permuted and sign-flipped per blob so that the blobs' subspaces differ) + a small
    isotropic term, rounded and clipped to [0, 255].  The low intrinsic dimension is what makes a true nearest
    neighbour distinguishable by a 16-byte code ("raise intra-blob structure").
    DEEP-shape (d=96 float, L2-normalised): the same construction in floats."""

    def __init__(self, w, torch, dev):
        self.w, self.torch, self.dev = w, torch, dev
        g = torch.Generator(device=dev)
        g.manual_seed(1234 * 7919 + 13)
        d = w["d"]
        self.nblob = nblob = max(w["nc"] // 4, 4)
        nsup = max(nblob // 8, 1)
        r = self.r = 12
        if w["kind"] == "sift":
            sup = (torch.empty(nsup, d, device=dev).exponential_(1.0, generator=g) +
                   torch.empty(nsup, d, device=dev).exponential_(1.0, generator=g)) * 15.0     # Gamma(2, 15)
            self.means = (sup[torch.randint(0, nsup, (nblob,), device=dev, generator=g)] +
                          torch.randn(nblob, d, device=dev, generator=g) * 12.0).clamp_(0, 255)
            self.s_lr, self.s_iso = 30.0, 3.0
        else:
            sup = torch.randn(nsup, d, device=dev, generator=g)
            self.means = sup[torch.randint(0, nsup, (nblob,), device=dev, generator=g)] + \
                torch.randn(nblob, d, device=dev, generator=g) * 0.5
            self.s_lr, self.s_iso = 0.8, 0.08
        F = torch.randn(r, d, device=dev, generator=g)
        self.F = F / F.norm(dim=1, keepdim=True)
        self.nbank = 64
        self.perm = torch.stack([torch.randperm(d, device=dev, generator=g) for _ in range(self.nbank)])
        self.sign = (torch.randint(0, 2, (self.nbank, d), device=dev, generator=g).float() * 2 - 1)

    def draw(self, n, seed):
        torch, dev, w = self.torch, self.dev, self.w
        g = torch.Generator(device=dev)
        g.manual_seed(seed)
        which = torch.randint(0, self.nblob, (n,), device=dev, generator=g)
        bank = which % self.nbank
        lr = (torch.randn(n, self.r, device=dev, generator=g) @ self.F) * self.s_lr
        lr = torch.gather(lr, 1, self.perm[bank]) * self.sign[bank]
        x = self.means[which] + lr + torch.randn(n, w["d"], device=dev, generator=g) * self.s_iso
        if w["kind"] == "sift":
            x = x.round_().clamp_(0, 255)
        else:
            x = x / x.norm(dim=1, keepdim=True)
        return x.contiguous()


def _gt_update(torch, w, queries, qn, x, b0, best_d, best_i):
    """Exact nearest base vector of every query (harness-side brute force for the recall figure).  Byte-valued data:
    TF32 products are exact (8-bit operands, fp32 accumulation of < 2^24 integers); DEEP: plain fp32."""
    exact_tf32 = w["kind"] == "sift"
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = bool(exact_tf32)
    try:
        for c0 in range(0, x.shape[0], 250_000):
            xc = x[c0:c0 + 250_000]
            dd = qn[:, None] - 2.0 * (queries @ xc.T) + (xc * xc).sum(1)[None]
            md, mi = dd.min(1)
            upd = md < best_d
            best_d.copy_(torch.where(upd, md, best_d))
            best_i.copy_(torch.where(upd, mi + (b0 + c0), best_i))
            del dd
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old


def cache_key(w):
    import hashlib
    h = hashlib.sha1(json.dumps(w, sort_keys=True).encode())
    h.update(b"gen%d" % GEN_VERSION)
    try:
        with open(pkg().LIB_PATH, "rb") as f:
            h.update(hashlib.sha1(f.read()).digest())       # code version: the built library itself
    except OSError:
        pass
    return h.hexdigest()[:16]


def cache_root():
    for base in (os.environ.get("IVFHNSW_BENCH_CACHE"), "/dev/shm", tempfile.gettempdir()):
        if base and os.path.isdir(base) and os.access(base, os.W_OK):
            return os.path.join(base, "ivfb200_cache")
    return None


def cache_dir_for(w):
    root = cache_root()
    if root is None or os.environ.get("IVFHNSW_BENCH_NO_CACHE") or w["nb"] < 20_000_000:
        return None
    return os.path.join(root, cache_key(w))


def save_cache(cdir, idx, w, cent_h, queries, gt0, log):
    """The built index as flat arrays (the vectors write() serialises), so that the two arms the driver runs back to
    back on one box -- and the cpu_baseline leg -- build a 1B index once."""
    try:
        os.makedirs(os.path.dirname(cdir), exist_ok=True)
        ix = idx.export_index_dict()
        need = sum(v.nbytes for v in ix.values() if isinstance(v, np.ndarray)) + cent_h.nbytes
        st = os.statvfs(os.path.dirname(cdir))
        if st.f_bavail * st.f_frsize < need * 1.2 + (4 << 30):
            log("index cache skipped: %.1f GB needed, %.1f GB free" % (need / 1e9, st.f_bavail * st.f_frsize / 1e9))
            return None
        tmp = cdir + ".tmp%d" % os.getpid()
        os.makedirs(tmp, exist_ok=True)
        for key, v in ix.items():
            if isinstance(v, np.ndarray):
                np.save(os.path.join(tmp, "ix_%s.npy" % key), v)
        np.save(os.path.join(tmp, "cent.npy"), cent_h)
        np.save(os.path.join(tmp, "pqc.npy"), idx.pq_centroids)
        np.save(os.path.join(tmp, "norm.npy"), idx.norm_centroids)
        if idx.opq_matrix is not None:
            np.save(os.path.join(tmp, "opq.npy"), idx.opq_matrix)
        np.save(os.path.join(tmp, "queries.npy"), queries.cpu().numpy())
        np.save(os.path.join(tmp, "gt0.npy"), gt0.cpu().numpy())
        with open(os.path.join(tmp, "done"), "w") as f:
            f.write("ok")
        if os.path.exists(cdir):
            import shutil
            shutil.rmtree(tmp, ignore_errors=True)
        else:
            os.rename(tmp, cdir)
        return ix
    except Exception as ex:       # a cache is an optimisation: never lose the run over it
        log("index cache not written: %r" % (ex,))
        return None


def load_cache(cdir, w, torch, dev, log):
    P = pkg()
    cls = P.IndexIVF_HNSW_Grouping if w["nsubc"] else P.IndexIVF_HNSW
    args = (w["d"], w["nc"], w["M"], 8) + ((w["nsubc"],) if w["nsubc"] else ())
    idx = cls(*args, device=dev.index or 0)
    # memory-mapped: the cache lives in tmpfs, so the arrays are used in place -- N ranks loading a 25 GB index must not
    # each copy it into their own address space (8 x 21 GB + the tmpfs itself would not fit the host)
    ld = lambda n: np.load(os.path.join(cdir, n), mmap_mode="r")
    cent_h = ld("cent.npy")
    idx.build_quantizer(centroids=cent_h)
    opq = ld("opq.npy") if os.path.exists(os.path.join(cdir, "opq.npy")) else None
    idx.set_pq(ld("pqc.npy"), ld("norm.npy"), opq)
    idx.do_opq = w["opq"]
    ix = {f[3:-4]: ld(f) for f in os.listdir(cdir) if f.startswith("ix_")}
    idx.load_index_dict(ix)
    if w["opq"]:
        idx.rotate_quantizer()
    idx.nprobe, idx.max_codes, idx.do_pruning = w["nprobe"], w["max_codes"], w["pruning"]
    queries = torch.from_numpy(np.array(ld("queries.npy"))).to(dev)
    gt0 = torch.from_numpy(np.array(ld("gt0.npy"))).to(dev)
    log("index loaded from cache %s (ntotal %d)" % (cdir, idx.ntotal))
    return idx, queries, gt0, cent_h


def build_index(w, torch, dev, log, dist=None, rank=0, world=1, sharded=False):
    """Everything on the GPU through the library's own entry points. Returns (index, queries_dev, gt0, centroids).

    IVFADC: one pass (assign + encode per 1M batch, IndexIVF_HNSW.cpp:75-138).
    Grouping: the reference's bounded passes (tests/test_ivfhnsw_grouping_sift1b.cpp:138-201): pass 1 assigns every
    batch (K-A1, the 2.7e17-FLOP part; split over the ranks when there are several, labels all-gathered), then the
    groups are taken range by range: the batches are re-drawn, the points of the range collected, add_points_dev +
    flush_groups encodes them and drops the raw vectors (peak HBM = one range, not the base set)."""
    sharded = bool(sharded and dist is not None and w["nsubc"] > 0)
    cdir = None if sharded else cache_dir_for(w)
    if cdir and os.path.exists(os.path.join(cdir, "done")):
        return load_cache(cdir, w, torch, dev, log)
    P = pkg()
    t0 = time.time()
    G = Gen(w, torch, dev)
    cls = P.IndexIVF_HNSW_Grouping if w["nsubc"] else P.IndexIVF_HNSW
    args = (w["d"], w["nc"], w["M"], 8) + ((w["nsubc"],) if w["nsubc"] else ())
    idx = cls(*args, device=dev.index or 0)
    # coarse centroids: Lloyd iterations driven by K-A1 itself (k-means is not on the path; any file works)
    learn = G.draw(w["nlearn"], 1234 + 2)
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
    del learn, cent
    torch.cuda.empty_cache()
    bs = 1_000_000
    nb = w["nb"]
    chunks = [(b0, min(bs, nb - b0)) for b0 in range(0, nb, bs)]
    seed_of = lambda b0: 99991 + b0
    queries = G.draw(w["nq"], 1234 + 1)
    if w.get("enn"):      # ENN: the queries ARE the first base vectors, ground truth = identity
        queries = G.draw(chunks[0][1], seed_of(0))[:w["nq"]].clone()
    ngt = min(w["nq"], w.get("gt_queries", w["nq"]))      # recall is measured on the first ngt queries
    qg = queries[:ngt]
    qn = (qg * qg).sum(1)
    best_d = torch.full((ngt,), float("inf"), device=dev)
    best_i = torch.zeros(ngt, dtype=torch.long, device=dev)
    grouping = w["nsubc"] > 0
    split = dist is not None and grouping and (nb >= 20_000_000 or sharded)     # pass 1 split over the ranks
    labels = torch.zeros(nb, dtype=torch.int32, device=dev) if grouping else None
    for ci, (b0, n) in enumerate(chunks):
        if split and ci % world != rank:
            continue
        x = G.draw(n, seed_of(b0))
        if grouping:
            idx.assign_dev(x, labels[b0:b0 + n])                       # K-A1: IndexIVF_HNSW::assign
        else:
            ids = torch.arange(b0, b0 + n, dtype=torch.int32, device=dev)
            idx.add_points_dev(x, ids, None)                            # assign + encode
        _gt_update(torch, w, qg, qn, x, b0, best_d, best_i)
        del x
        if ci % 100 == 99:
            log("pass 1: %d / %d batches, %.1fs" % (ci + 1, len(chunks), time.time() - t0))
    if split:
        dist.all_reduce(labels, op=dist.ReduceOp.SUM)          # disjoint slices: the sum is the concatenation
        allb = [torch.empty_like(best_d) for _ in range(world)]
        alli = [torch.empty_like(best_i) for _ in range(world)]
        dist.all_gather(allb, best_d)
        dist.all_gather(alli, best_i)
        sb, si = torch.stack(allb), torch.stack(alli)
        win = sb.argmin(0)
        best_i = si.gather(0, win[None])[0]
    if grouping:
        idx.sync()                                             # also checks the coarse stage's overflow flags
        log("pass 1 (assign) done %.1fs" % (time.time() - t0))
        # group ranges of at most ~pass_points points each
        per_pass = int(w.get("pass_points", 48_000_000))
        cnt = torch.zeros(w["nc"], dtype=torch.long, device=dev)
        for b0, n in chunks:
            cnt += torch.bincount(labels[b0:b0 + n].long(), minlength=w["nc"])
        cum = torch.cumsum(cnt, 0).cpu().numpy()
        own_lo, own_hi = 0, w["nc"]
        if sharded:
            # SHARDED CONSTRUCTION: this rank keeps and encodes only the groups it owns (1/world of the points, of the
            # encode work and of the HBM); the per-list tables are summed over the ranks at the end
            sharding = importlib.import_module("ivf-hnsw_b200.sharding")
            own_lo, own_hi = sharding.shard_range(rank, world, w["nc"])
            idx.set_shard(own_lo, own_hi)
        bounds, lo = [], own_lo
        while lo < own_hi:
            base = cum[lo - 1] if lo else 0
            hi = int(np.searchsorted(cum, base + per_pass, side="right"))
            hi = max(hi, lo + 1)
            bounds.append((lo, min(hi, own_hi)))
            lo = hi
        for pi, (lo, hi) in enumerate(bounds):
            for b0, n in (chunks if (len(bounds) > 1 or sharded) else []):
                lab = labels[b0:b0 + n]
                sel = ((lab >= lo) & (lab < hi)).nonzero().squeeze(1)
                if sel.numel() == 0:
                    continue
                x = G.draw(n, seed_of(b0))
                idx.add_points_dev(x[sel].contiguous(), (sel + b0).int().contiguous(), lab[sel].contiguous())
                del x
            if len(bounds) == 1 and not sharded:      # small index: one range, batches added as they are drawn
                for b0, n in chunks:
                    x = G.draw(n, seed_of(b0))
                    idx.add_points_dev(x, torch.arange(b0, b0 + n, dtype=torch.int32, device=dev), labels[b0:b0 + n])
                    del x
            idx.flush_groups()
            log("pass 2: group range %d / %d encoded, %.1fs" % (pi + 1, len(bounds), time.time() - t0))
        del labels
    idx.compute_centroid_norms()       # also the flush point (finalize)
    if w["nsubc"]:
        idx.compute_inter_centroid_dists()
    if w["opq"]:
        idx.rotate_quantizer()
    if sharded:
        sharding.exchange_list_meta(torch, dist, idx, dev)
    torch.cuda.empty_cache()
    log("index built %.1fs (ntotal %d)" % (time.time() - t0, idx.ntotal))
    idx.nprobe, idx.max_codes, idx.do_pruning = w["nprobe"], w["max_codes"], w["pruning"]
    if cdir and rank == 0:
        save_cache(cdir, idx, w, cent_h, queries, best_i, log)
        log("index cached %.1fs" % (time.time() - t0))
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
def reference_workdir(idx, w, cent_h, log):
    """Everything the compiled reference needs, in its own file formats: x.index (written natively by the library,
    byte-compatible), pq/norm/opq files, the UNROTATED centroid file.  Kept next to the index cache so that the
    reference arm and the cpu_baseline leg of the other arm share the export and the HNSW graph."""
    P = pkg()
    cdir = cache_dir_for(w)
    workdir = (cdir + "_ref") if cdir else tempfile.mkdtemp(prefix="ivfb200_ref_")
    if os.path.exists(os.path.join(workdir, "exported")):
        return workdir
    os.makedirs(workdir, exist_ok=True)
    t0 = time.time()
    idx.write(os.path.join(workdir, "x.index"))
    P.write_ProductQuantizer(os.path.join(workdir, "pq.pq"), w["d"], w["M"], idx.pq_centroids)
    P.write_ProductQuantizer(os.path.join(workdir, "norm.pq"), 1, 1, idx.norm_centroids)
    if w["opq"]:
        P.write_VectorTransform(os.path.join(workdir, "opq.mat"), idx.opq_matrix)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as O
    O.write_fvecs(os.path.join(workdir, "centroids.fvecs"), cent_h)   # UNROTATED file, like the reference's input
    with open(os.path.join(workdir, "exported"), "w") as f:
        f.write("ok")
    log("reference files written %.1fs" % (time.time() - t0))
    return workdir


def run_reference_cpu(w, workdir, queries_h, gt0, nq_sample, nproc, steps=1, warmup=0, log=None):
    """Times IndexIVF_HNSW[_Grouping]::search of the compiled reference (trace-stripped build) on the host
    cores: the index is read ONCE, then `nproc` forked workers each search a disjoint query slice sequentially,
    exactly like tests/test_ivfhnsw_sift1b.cpp:192-213.  Returns the result dict of
    oracle/ref_worker.py; <workdir>/parity.npz holds the reference's answers and HNSW candidate lists."""
    np.save(os.path.join(workdir, "queries.npy"), queries_h[:nq_sample])
    np.save(os.path.join(workdir, "gt0.npy"), gt0[:nq_sample])
    cfg = dict(w, workdir=workdir, nproc=nproc, nq_sample=nq_sample, steps=steps, warmup=warmup, k_recall=100)
    with open(os.path.join(workdir, "cfg.json"), "w") as f:
        json.dump(cfg, f)
    env = dict(os.environ, OMP_NUM_THREADS="1", OPENBLAS_NUM_THREADS="1")
    t0 = time.time()
    subprocess.check_call([sys.executable, os.path.join(ROOT, "oracle", "ref_worker.py"), workdir, "run"], env=env,
                          stdout=subprocess.DEVNULL)
    res = json.load(open(os.path.join(workdir, "result.json")))
    if log:
        log("reference: HNSW %.0fs, read %.0fs, total %.0fs" % (res["hnsw_s"], res["read_s"], time.time() - t0))
    return res


def parity_against_reference(idx, w, workdir, queries_h, k, gt0_h=None):
    """The north_star bar, checked on the BENCHED index inside the run: the reference's own coarse candidate lists
    (its HNSW's answers, dumped by the CPU leg) go through ivfhnsw_b200_search_given_coarse; labels must be identical
    up to equal-distance ties and distances within 1e-4 relative."""
    z = np.load(os.path.join(workdir, "parity.npz"))
    Dr, Lr, cd, ci = z["D"], z["L"], z["cd"], z["ci"]
    n = Dr.shape[0]
    D, L = idx.search2(k, queries_h[:n], cd, ci)
    o = np.argsort(Dr, axis=1, kind="stable")
    Dr, Lr = np.take_along_axis(Dr, o, 1), np.take_along_axis(Lr, o, 1)
    fin = np.abs(Dr) < 1e30
    # the terms that cancel in an ADC distance are O(||c||^2): same guard as the parity tests (tests/conftest.py)
    scale = float(idx.get_centroid_norms().max())
    same_fin = bool(np.array_equal(fin, np.abs(D) < 1e30))
    rel = np.abs(D[fin] - Dr[fin]) / np.maximum(np.abs(Dr[fin]), 1e-2 * scale) if same_fin and fin.any() else np.array([np.inf])
    bad = 0
    for i in range(n):
        if set(L[i].tolist()) == set(Lr[i].tolist()):
            continue
        kth = Dr[i][-1]
        tol = 1e-4 * max(abs(float(kth)), 1e-2 * scale)
        a = {int(l) for l, dd in zip(L[i], D[i]) if dd < kth - tol}
        b = {int(l) for l, dd in zip(Lr[i], Dr[i]) if dd < kth - tol}
        bad += a != b
    ours_r10 = None
    if gt0_h is not None:      # our recall on the reference's own sample and candidate lists
        ours_r10 = float(np.mean([gt0_h[i] in L[i] for i in range(n)]))
    return {"queries": int(n), "labels_equal_up_to_ties": bad == 0, "queries_with_label_mismatch": int(bad),
            "recall_at_10_on_reference_candidates": ours_r10,
            "labels_identical_rows": int(sum(np.array_equal(L[i], Lr[i]) for i in range(n))),
            "max_rel_dist_err": float(rel.max()), "tolerance": 1e-4,
            "how": "reference's HNSW candidate lists -> ivfhnsw_b200_search_given_coarse on the benched index"}


# --------------------------------------------------------------------------- assign workload (configs[2])
def run_assign(args, w, torch, dev, dist, rank, world, local, log, emit):
    """IndexIVF_HNSW::assign (IndexIVF_HNSW.cpp:68-72) of one batch of base vectors per step: K-A1 = the coarse
    GEMM + exact re-rank with nprobe 1.  value: device-resident float batch; e2e: the C-ABI assign_u8 call with a
    HOST uint8 batch (what mkindex streams from the .bvecs file), labels read back."""
    P = pkg()
    t0 = time.time()
    G = Gen(w, torch, dev)
    cent = G.draw(w["nc"], 4242)          # any centroid file works for the timing
    if dist is not None:
        dist.broadcast(cent, src=0)
    idx = P.IndexIVF_HNSW(w["d"], w["nc"], w["M"], 8, device=dev.index or 0)
    if args.coarse_impl >= 0:
        idx.set_option("coarse_impl", args.coarse_impl)
    idx.build_quantizer(centroids=cent.cpu().numpy())
    n = w["nb"]
    x = G.draw(n, 99991 + rank)
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
                           "algorithmic": "2*nc*d FLOP per vector (coarse_impl 3 issues exactly that: ONE fp16 product, the dropped cross terms are covered by the select margin + exact fp32 re-rank)", "traffic": None,
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
    ap.add_argument("--workload", default=os.environ.get("IVFHNSW_BENCH_WORKLOAD", DEFAULT_WORKLOAD))
    ap.add_argument("--cpu-sample", type=int, default=0, help="queries for the CPU baseline (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--list-shards", type=int, default=int(os.environ.get("IVFHNSW_BENCH_LIST_SHARDS", "0")),
                    help="list shards per query group (0 = the smallest count whose lists fit one GPU)")
    ap.add_argument("--coarse-impl", type=int, default=int(os.environ.get("IVFHNSW_COARSE_IMPL", "-1")))
    ap.add_argument("--sharded-build", action="store_true",
                    help="N > 1, Grouping workloads: every rank ingests and keeps only the lists it owns "
                         "(ivfhnsw_b200_set_shard before the build + exchange of the per-list tables); measures the "
                         "list-sharded layout alone")
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
    t_build = time.time()
    sharded_build = bool(args.sharded_build and world > 1 and w["nsubc"] > 0 and args.impl == "ours")
    idx, queries, gt0, cent_h = build_index(w, torch, dev, log, dist, rank, world, sharded=sharded_build)
    t_build = time.time() - t_build
    if args.coarse_impl >= 0:
        idx.set_option("coarse_impl", args.coarse_impl)
    nq, k, nprobe = w["nq"], w["k"], w["nprobe"]
    ngt = int(gt0.shape[0])                      # recall is measured on the first ngt queries
    ncores = len(os.sched_getaffinity(0))
    workload_name = {"c4": "SIFT1B-shape 1Bx128 (byte-valued), 2^20 centroids, 16-byte OPQ, Grouping(64)+Pruning, nprobe=128, "
                           "max_codes=100000, k=10, 10k queries/step (BASELINE configs[3], the configuration the metric is "
                           "quoted on)",
                     "c2": "DEEP-shape 10Mx96 f32, 2^16 centroids, 16-byte PQ, Grouping(64)+Pruning, nprobe=64, "
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
                     "c5b": "SIFT1B ENN variant at 1/50 scale with 8-byte (64-bit) codes: 20Mx128, 2^15 centroids, IVFADC, "
                            "nprobe=1, k=1, queries = first 10k base vectors (BASELINE configs[4] scaled)",
                     "tiny": "tiny development workload"}[args.workload]
    config = {"workload": workload_name, "nb": w["nb"], "d": w["d"], "nc": w["nc"], "code_size": w["M"],
              "nsubc": w["nsubc"], "pruning": w["pruning"], "opq": w["opq"], "nprobe": nprobe, "max_codes": w["max_codes"],
              "k": k, "nq_per_step": nq, "l2": "flushed between timed steps (256 MiB write)",
              "sharding": "lists by centroid range, NCCL exchange of work records and partial top-k" if world > 1 else "none"}

    def recalls(Ls):
        """Recall@R as the reference defines it (tests/test_ivfhnsw_sift1b.cpp:175-176,203-207): the true nearest
        neighbour is among the first R returned labels."""
        hit = Ls[:ngt] == gt0[:, None]
        return {"r1": float(hit[:, :1].any(1).float().mean().item()),
                "r10": float(hit[:, :10].any(1).float().mean().item()),
                "r100": float(hit[:, :100].any(1).float().mean().item())}

    # ------------------------------------------------------------------ reference arm
    if args.impl == "reference":
        workdir = reference_workdir(idx, w, cent_h, log)
        idx.close()
        torch.cuda.empty_cache()
        nproc = ncores
        sample = args.cpu_sample or min(ngt, max(nproc * 40, 400))
        res = run_reference_cpu(w, workdir, queries.cpu().numpy(), gt0.cpu().numpy(), sample, nproc,
                                steps=args.steps, warmup=args.warmup, log=log)
        line = {"impl": "reference", "metric": "QPS at fixed Recall@10 (k=10)", "value": res["qps"], "unit": "queries/s",
                "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": res["ms_per_step"],
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": config, "recall_at_10": res["recall"]["r10"], "recall": res["recall"],
                "cpu_baseline": {"value": res["qps"], "unit": "queries/s", "cores": nproc, "kind": "reference",
                                 "sample": "%d queries of the same batch, %d single-thread processes (forked after one "
                                           "read()) of the compiled reference over the GPU-built .index, efSearch %d "
                                           "(%.0f us/query/thread)" % (sample, nproc, w["efs"], res["us_per_query"])},
                "e2e": {"value": res["qps"], "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        emit(line)
        return 0

    # ------------------------------------------------------------------ our arm
    # N > 1: the inverted lists are sharded over ALL ranks by centroid range (north_star).  The
    # headline is WEAK scaling of that layout: every rank brings its own batch of nq queries (prepared at home: coarse
    # stage + pruning pass), the work records are exchanged, every rank scans the lists it owns for the queries of all
    # ranks, partial top-k lists are exchanged and merged.  The same layout on ONE fixed batch (strong scaling) and the
    # replica layout (every rank searching the whole index alone) are measured beside it.
    stream = torch.cuda.Stream(device=dev)   # a real (non-default) stream shared by torch and the library
    torch.cuda.set_stream(stream)
    idx.set_stream(stream.cuda_stream)
    D = torch.empty(nq, k, dtype=torch.float32, device=dev)
    L = torch.empty(nq, k, dtype=torch.int64, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def single_search(q_dev):
        idx.search_dev(k, q_dev, D, L)

    def make_sharded(nq_home):
        """Sharded search of world * nq_home queries per step of which this rank is the home of rows [rank * nq_home,
        (rank + 1) * nq_home): sharding.PackedShardedSearch (all-to-all-v of packed per-owner records)."""
        ps = sharding.PackedShardedSearch(torch, dist, idx, rank, world, nq_home, k, dev)
        return ps.search, ps

    def timed_run(search, q_dev, steps, warmup, clocks):
        for _ in range(max(warmup, 3)):
            search(q_dev)
        torch.cuda.synchronize()
        idx.reset_stats()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        torch.cuda.nvtx.range_push("timed")     # ncu --nvtx --nvtx-include "timed/" profiles exactly this region
        with clocks:
            for s_, e_ in evs:
                flush.zero_()
                s_.record()
                search(q_dev)
                e_.record()
            torch.cuda.synchronize()
        torch.cuda.nvtx.range_pop()
        if dist is not None:
            dist.barrier()
        ms = sum(s_.elapsed_time(e_) for s_, e_ in evs)
        st = idx.stats()
        if dist is not None:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, st

    clk = Clocks(local)
    sharded_ok, layouts = None, None
    if sharded_build:
        # no rank holds the whole index: only the list-sharded layout exists.  Its answers are checked through recall
        # (the index is bit-identical to the one a single GPU builds, tests/test_build_gpu.py::
        # test_sharded_construction_equals_one_index, so recall must equal the N=1 line's)
        free_b, total_b = torch.cuda.mem_get_info()
        q_weak = torch.cat([torch.roll(queries, shifts=-997 * r, dims=0) for r in range(world)])
        w_search, w_ps = make_sharded(nq)
        ms, st = timed_run(w_search, q_weak, args.steps, args.warmup, clk)
        Lh = w_ps.L.clone()
        hit = Lh[:ngt] == gt0[:, None]
        line = {"metric": "QPS at fixed Recall@10 (k=10)", "mode": "sharded_build",
                "value": nq * world * args.steps / (ms / 1e3), "unit": "queries/s", "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "u8 codes, f32 accumulate", "data": "synthetic",
                "config": dict(config, nq_per_step=nq * world,
                               sharding="lists sharded over %d ranks at CONSTRUCTION: every rank assigned 1/%d of the base "
                                        "vectors (labels all-reduced), then ingested, encoded and keeps only the groups of "
                                        "its centroid range; per-list tables summed by all-reduce" % (world, world)),
                "build_seconds": t_build, "hbm_used_bytes_after_build_rank0": int(total_b - free_b),
                "recall_rank0": {"r1": float(hit[:, :1].any(1).float().mean().item()),
                                 "r10": float(hit[:, :10].any(1).float().mean().item())},
                "record_exchange_MB_received_per_rank_per_step": w_ps.last_recv_words * 4 / 1e6,
                "clocks": clk.summary(), "gpu_launches": int(st.kernel_launches)}
        if rank == 0:
            emit(line)
        idx.close()
        dist.destroy_process_group()
        return 0
    if world == 1:
        idx.set_shard(0, w["nc"])
        ms, st = timed_run(single_search, queries, args.steps, args.warmup, clk)
        value = nq * args.steps / (ms / 1e3)
        scaling = "weak"
        search = single_search
        q_timed = queries
    else:
        # replicas first (also gives the unsharded answer for the equality check)
        idx.set_shard(0, w["nc"])
        ms_rep, _ = timed_run(single_search, queries, args.steps, args.warmup, contextlib.nullcontext())
        D_ref, L_ref = D.clone(), L.clone()
        idx.set_shard(*sharding.shard_range(rank, world, w["nc"]))
        # strong: ONE batch of nq queries over all ranks
        per = (nq + world - 1) // world
        nq_s = per * world
        q_strong = torch.cat([queries, queries[:nq_s - nq]]) if nq_s > nq else queries
        s_search, s_ps = make_sharded(per)
        ms_str, _ = timed_run(s_search, q_strong, args.steps, args.warmup, contextlib.nullcontext())
        lo_s, hi_s = rank * per, min((rank + 1) * per, nq)
        ok_here = bool(torch.equal(s_ps.L[:hi_s - lo_s], L_ref[lo_s:hi_s]) and torch.equal(s_ps.D[:hi_s - lo_s], D_ref[lo_s:hi_s]))
        # weak (headline): every rank brings its own nq queries (rank r's batch = the same seeded batch rotated by
        # r * 997 rows, so that all ranks hold the full tensor without a broadcast in the timed region)
        q_weak = torch.cat([torch.roll(queries, shifts=-997 * r, dims=0) for r in range(world)])
        w_search, w_ps = make_sharded(nq)
        ms, st = timed_run(w_search, q_weak, args.steps, args.warmup, clk)
        value = nq * world * args.steps / (ms / 1e3)
        scaling = "weak"
        search, q_timed = w_search, q_weak
        # a rank's home batch is the seeded batch rotated by 997 * rank rows: its answer must be the single-GPU answer
        # rotated the same way, bit for bit
        ok_here = ok_here and bool(torch.equal(w_ps.L, torch.roll(L_ref, shifts=-997 * rank, dims=0)) and
                                   torch.equal(w_ps.D, torch.roll(D_ref, shifts=-997 * rank, dims=0)))
        okt = torch.tensor([1 if ok_here else 0], device=dev)
        dist.all_reduce(okt, op=dist.ReduceOp.MIN)
        sharded_ok = bool(okt.item())
        if rank == 0:
            D.copy_(w_ps.D); L.copy_(w_ps.L)
        exchange_mb = w_ps.last_recv_words * 4 / 1e6
        layouts = {"list_sharded_weak": {"value": value, "ms_per_step": ms / args.steps, "nq_per_step": nq * world},
                   "list_sharded_strong": {"value": nq * args.steps / (ms_str / 1e3), "ms_per_step": ms_str / args.steps,
                                           "nq_per_step": nq},
                   "replicas": {"value": nq * world * args.steps / (ms_rep / 1e3), "ms_per_step": ms_rep / args.steps,
                                "nq_per_step": nq * world, "note": "every rank searches the whole index alone, no collective"},
                   "record_exchange_MB_received_per_rank_per_step": exchange_mb}
        config["nq_per_step"] = nq * world
        config["sharding"] = ("lists sharded over all %d ranks by centroid range; per step every rank prepares its own %d "
                              "queries, packs their work records per owning rank, ONE NCCL all-to-all-v moves them, every "
                              "rank scans its lists for all %d queries, partial top-k go home by all-to-all and are merged"
                              % (world, nq, nq * world))
        config["layouts"] = layouts
        config["sharded_equals_single_gpu"] = sharded_ok
    launches = int(st.kernel_launches)
    scanned = int(st.scanned_codes)
    ms_per_step = ms / args.steps

    # recall@1/10/100 of the benched index (one untimed k=100 search of the first ngt queries)
    kr = 100
    Dr = torch.empty(ngt, kr, dtype=torch.float32, device=dev)
    Lr = torch.empty(ngt, kr, dtype=torch.int64, device=dev)
    idx.set_shard(0, w["nc"])
    idx.search_dev(kr, queries[:ngt].contiguous(), Dr, Lr)
    torch.cuda.synchronize()
    recall = recalls(Lr)
    if world > 1:
        idx.set_shard(*sharding.shard_range(rank, world, w["nc"]))

    # per-stage device times (separate pass: the timing mode adds one sync per stage)
    idx.reset_stats()
    idx.set_timing(True)
    for _ in range(3):
        flush.zero_()
        search(q_timed)
    torch.cuda.synchronize()
    st2 = idx.stats()
    idx.set_timing(False)
    n_t = max(1, int(st2.scan_launches))
    n_p = max(1, int(st2.prep_launches))
    scan_ms, coarse_ms, prep_ms = st2.scan_ms / n_t, st2.coarse_ms / n_p, st2.prep_ms / n_p
    gemm_ms, g1_ms, lut_ms = st2.gemm_ms / n_p, st2.g1_ms / n_p, st2.lut_ms / n_t
    scan_kernel_ms = scan_ms - lut_ms
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
    # ~1/world of them.  Per LAUNCH: the timing pass ran 3 steps of n_t / 3 scan launches each.
    # (N > 1: the counter holds the codes of the queries this rank PREPARED, over all shards; by symmetry that is the
    # number of codes this rank scans for the queries of all ranks)
    scan_bytes = st2.scanned_codes / float(n_t) * code_bytes
    scan_gbs = scan_bytes / (scan_kernel_ms * 1e-3) / 1e9 if scan_kernel_ms > 0 else 0.0
    nq_prep = nq                                   # queries this rank prepares per step
    gemm_flop = 2.0 * w["nc"] * w["d"] * nq_prep * 3 / n_p
    coarse_tf = gemm_flop / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else 0.0
    # K-G1b: d*4 bytes per DISTINCT neighbour-centroid row gathered
    g1_bytes = (st2.g1_rows / float(n_p)) * w["d"] * 4
    g1_gbs = g1_bytes / (g1_ms * 1e-3) / 1e9 if g1_ms > 0 else 0.0
    nq_step = nq * world if world > 1 else nq
    stages = {"coarse_ms": coarse_ms * n_p / 3, "prep_ms": prep_ms * n_p / 3, "scan_ms": scan_ms * n_t / 3,
              "gemm_kernel_ms": gemm_ms * n_p / 3, "g1b_kernel_ms": g1_ms * n_p / 3, "lut_kernel_ms": lut_ms * n_t / 3,
              "scan_kernel_ms": scan_kernel_ms * n_t / 3, "launches_per_step": {"prep": n_p / 3, "scan": n_t / 3},
              "scan_GBps_algorithmic": scan_gbs, "coarse_TFLOPs_algorithmic": coarse_tf,
              "g1_distinct_rows_per_query": st2.g1_rows / 3.0 / max(1, nq_prep),
              "scanned_codes_per_query": st2.scanned_codes / 3.0 / nq_step,
              "walked_codes_per_query": st2.walked_codes / 3.0 / nq_step}
    # per-launch DRAM traffic of each kernel: measured by `ncu --set full` on THIS code (profiles/ncu_traffic.py writes
    # the file with a hash of csrc/); a file captured from other kernel sources is refused
    traffic, traffic_note = {}, "no capture for this workload"
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "r02_ncu_traffic.json")))
        if tj.get("csrc_sha1") == csrc_hash():
            traffic = tj.get(args.workload, {})
            traffic_note = "ncu --set full capture of these kernel sources (csrc sha1 %s)" % tj.get("csrc_sha1", "")[:12]
        else:
            traffic_note = "stale capture refused (taken from other kernel sources)"
    except Exception:
        pass
    rl = {
        "scan_kernel": {"kernel": "scan_kernel (K-S1/K-S2)", "bound": "hbm", "achieved": scan_gbs, "peak": hbm_peak,
                        "unit": "GB/s", "frac": scan_gbs / hbm_peak, "ms": scan_kernel_ms,
                        "algorithmic": "%d B per code of the sub-lists the reference rule scans, per launch" % code_bytes,
                        "traffic": traffic.get("scan_kernel") if world == 1 else None},
        "group_g1b_kernel": {"kernel": "group_g1b_kernel (K-G1b)", "bound": "hbm", "achieved": g1_gbs, "peak": hbm_peak,
                             "unit": "GB/s", "frac": g1_gbs / hbm_peak, "ms": g1_ms,
                             "algorithmic": "d*4 B per distinct neighbour-centroid row gathered (rows are L2 hits when "
                                            "the centroid table fits the 126 MB L2: %d MB here)" %
                                            (w["nc"] * w["d"] * 4 >> 20),
                             "traffic": traffic.get("group_g1b_kernel") if world == 1 else None},
        "coarse_tc_kernel": {"kernel": "coarse_tc_kernel (K-A2 GEMM)", "bound": "tensor", "achieved": coarse_tf,
                             "peak": tf_peak, "unit": "TFLOP/s", "frac": coarse_tf / tf_peak, "ms": gemm_ms,
                             "algorithmic": "2*nc*d FLOP per query (coarse_impl 3 issues exactly that: ONE fp16 product, the dropped cross terms are covered by the select margin + exact fp32 re-rank)",
                             "traffic": traffic.get("coarse_tc_kernel") if world == 1 else None},
    }
    # headline roofline: the larger (by live-measured time per step) of the two kernels whose roof is one of the two bound
    # types of the contract -- the ADC scan (HBM) and the coarse GEMM (tensor).  K-G1b stays under "others": its rows are
    # L2 hits whenever the centroid table fits L2, so its GB/s is not an HBM fraction (it can exceed the HBM peak).
    dom = max((rl["scan_kernel"], rl["coarse_tc_kernel"]),
              key=lambda r_: r_["ms"] * (n_t if r_ is rl["scan_kernel"] else n_p))
    rl["group_g1b_kernel"]["note"] = "rows are L2 hits when the centroid table fits L2: not an HBM fraction"
    roofline = dict(dom)
    roofline["peak_source"] = peak_src
    roofline["traffic_source"] = traffic_note
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
        # the host-buffer path must return exactly what the device-resident path returned
        e2e["equals_device_path"] = bool(torch.equal(Lh.to(dev), L) and torch.equal(Dh.to(dev), D))
        if not e2e["equals_device_path"]:
            log("ERROR: ivfhnsw_b200_search (host buffers) and search_dev disagree")
    else:
        # multi-GPU e2e: every rank copies ITS batch in from pinned host memory, the batches are all-gathered, the
        # sharded search runs, the rank reads its own rows of the answer back
        nqt = q_timed.shape[0]
        qh = q_timed[rank * nq:(rank + 1) * nq].cpu().pin_memory()
        Dh = torch.empty(nq, k, dtype=torch.float32).pin_memory()
        Lh = torch.empty(nq, k, dtype=torch.int64).pin_memory()
        qd = torch.empty_like(q_timed)

        def e2e_step():
            qd[rank * nq:(rank + 1) * nq].copy_(qh, non_blocking=True)
            dist.all_gather_into_tensor(qd, qd[rank * nq:(rank + 1) * nq])
            search(qd)
            Dh.copy_(w_ps.D, non_blocking=True)
            Lh.copy_(w_ps.L, non_blocking=True)
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
        e2e = {"value": nqt * args.steps / float(t.item()), "unit": "queries/s",
               "h2d_bytes_per_step": int(qh.numel() * 4) * world,
               "d2h_bytes_per_step": int(Dh.numel() * 4 + Lh.numel() * 8) * world}

    cpu_baseline, parity, ref_recall = None, None, None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            workdir = reference_workdir(idx, w, cent_h, log)
            sample = args.cpu_sample or min(ngt, max(ncores * 40, 400))
            qh_np = queries.cpu().numpy()
            # one untimed pass first: the forked workers touch the index pages (25 GB at C4) for the first time in it
            res = run_reference_cpu(w, workdir, qh_np, gt0.cpu().numpy(), sample, ncores, steps=3, warmup=1, log=log)
            ref_recall = res["recall"]
            cpu_baseline = {"value": res["qps"], "unit": "queries/s", "cores": ncores, "kind": "reference",
                            "sample": "%d queries of the same batch; %d single-thread processes (forked after one read()) "
                                      "of the compiled reference (trace-stripped Grouping) over the GPU-built .index, "
                                      "efSearch %d; %.0f us/query/thread; reference recall@10 %.3f" %
                                      (sample, ncores, w["efs"], res["us_per_query"], res["recall"]["r10"])}
            parity = parity_against_reference(idx, w, workdir, qh_np, k, gt0.cpu().numpy())
            Ls = L[:res["nq"]]          # full GPU path (exact coarse stage) on the reference's sample
            parity["recall_at_10_same_queries"] = {"ours": float((Ls == gt0[:res["nq"], None]).any(1).float().mean().item()),
                                                   "reference": res["recall"]["r10"]}
        except Exception as ex:   # the baseline is a reported extra; never lose the GPU line over it
            cpu_baseline = {"value": None, "unit": "queries/s", "cores": ncores, "kind": "reference",
                            "sample": "failed: %r" % (ex,)}
        if parity is not None and not parity["labels_equal_up_to_ties"]:
            log("PARITY FAILURE against the reference: %r" % (parity,))

    if rank == 0:
        line = {"metric": "QPS at fixed Recall@10 (k=10)", "value": value, "unit": "queries/s", "n_gpus": world,
                "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
                "scaling": scaling, "vs_baseline": None, "dtype": "f32",
                "data": "synthetic", "config": config, "recall_at_10": recall["r10"], "recall": recall,
                "reference_recall": ref_recall, "recall_queries": ngt, "parity": parity,
                "clocks": clk.summary(), "e2e": e2e, "gpu_launches": launches,
                "roofline": roofline, "stages": stages, "cpu_baseline": cpu_baseline,
                "scanned_codes_per_step": scanned // max(1, args.steps)}
        emit(line)
    if dist is not None:
        dist.destroy_process_group()
    return 0


def csrc_hash():
    """sha1 over the kernel sources: ties profiles/r02_ncu_traffic.json to the code it was captured from."""
    import hashlib
    h = hashlib.sha1()
    cs = os.path.join(ROOT, "ivf-hnsw_b200", "csrc")
    for f in sorted(os.listdir(cs)):
        if f.endswith((".cu", ".cuh", ".h", ".cpp")):
            with open(os.path.join(cs, f), "rb") as fh:
                h.update(f.encode() + b"\0" + fh.read())
    return h.hexdigest()


if __name__ == "__main__":
    sys.exit(main())
