"""TEST INFRASTRUCTURE: numpy statement of ivfhnsw_b200_pack_records_dev (include/ivfhnsw_b200.h) -- the split of the
fixed-stride per-query records into one packed stream per owning rank.  The CUDA kernels are compared with it on the
GPU; the gloo test uses it as the stand-in of the device step."""
import numpy as np


def record_words(nprobe, nsubc):
    return (5 * nprobe + nprobe * nsubc + 3) & ~3


def pack_records(rec, nprobe, nsubc, nc, nparts):
    """rec: uint32 [nq][stride].  -> (streams: list of uint32 arrays, cnt [nparts][nq], off [nparts][nq], pwords [nparts])"""
    nq = rec.shape[0]
    per = (nc + nparts - 1) // nparts
    W = 5 + nsubc
    cnt = np.zeros((nparts, nq), np.uint32)
    off = np.zeros((nparts, nq), np.uint32)
    chunks = [[] for _ in range(nparts)]
    for q in range(nq):
        r = rec[q]
        n, cid = r[nprobe:2 * nprobe], r[4 * nprobe:5 * nprobe]
        part = np.minimum(cid // per, nparts - 1)
        sub = r[5 * nprobe:5 * nprobe + nprobe * nsubc].reshape(nprobe, nsubc) if nsubc else None
        for p in range(nparts):
            sel = np.nonzero((n != 0) & (part == p))[0]
            cnt[p, q] = sel.size
            off[p, q] = sum(c.size for c in chunks[p])
            fields = [r[f * nprobe:(f + 1) * nprobe][sel] for f in range(5)]
            if nsubc:
                fields.append(sub[sel].ravel())
            chunks[p].append(np.concatenate(fields).astype(np.uint32) if sel.size else np.zeros(0, np.uint32))
    streams = [np.concatenate(c) if c else np.zeros(0, np.uint32) for c in chunks]
    pwords = np.array([s.size for s in streams], np.uint64)
    assert all(int(cnt[p].sum()) * W == streams[p].size for p in range(nparts))
    return streams, cnt, off, pwords
