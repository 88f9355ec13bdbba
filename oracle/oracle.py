"""TEST INFRASTRUCTURE ONLY -- ctypes loaders for the two CPU checkers.

* ``Port``  : oracle/libivfhnsw_port.so, the plain-C restatement (oracle/ivfhnsw_port.c).
* ``Ref``   : oracle/_ref/libivfhnsw_ref.so, the *reference itself* (compiled from
              /root/reference by oracle/Makefile) behind oracle/ref_shim.cpp.

Only tests/, __graft_entry__.smoke() and bench.py's cpu legs import this module.
The product package (ivf-hnsw_b200) never does.
"""
import ctypes as C
import os
import struct
import subprocess
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "libivfhnsw_port.so")
REF_SO = os.path.join(HERE, "_ref", "libivfhnsw_ref.so")
REF_SHIPPED_SO = os.path.join(HERE, "_ref", "libivfhnsw_ref_shipped.so")

f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
u32p = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")
u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")
i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
sz = C.c_size_t


def build(port=True, ref=True):
    """Compile the checkers (oracle/Makefile). ref is skipped where /root/reference is absent."""
    targets = (["port"] if port else []) + (["ref"] if ref and os.path.exists("/root/reference") else [])
    if targets:
        subprocess.check_call(["make", "-s", "-C", HERE, "-j8"] + targets)


def have_ref():
    return os.path.exists(REF_SO)


# --------------------------------------------------------------- file formats
def write_fvecs(path, x):
    x = np.ascontiguousarray(x, dtype=np.float32)
    n, d = x.shape
    out = np.empty((n, d + 1), dtype=np.float32)
    out[:, 0] = np.array([d], dtype=np.uint32).view(np.float32)[0]
    out[:, 1:] = x
    out.tofile(path)


def write_pq(path, d, M, centroids):
    """faiss write_ProductQuantizer layout (faiss/index_io.cpp:200-205): size_t d,M,nbits; size_t n; n*f32."""
    c = np.ascontiguousarray(centroids, dtype=np.float32).ravel()
    with open(path, "wb") as f:
        f.write(struct.pack("<QQQQ", d, M, 8, c.size))
        f.write(c.tobytes())


def write_opq(path, A):
    """faiss write_VectorTransform, generic LinearTransform (faiss/index_io.cpp:157-198)."""
    A = np.ascontiguousarray(A, dtype=np.float32)
    d = A.shape[0]
    with open(path, "wb") as f:
        f.write(b"LTra")
        f.write(struct.pack("<?", False))
        f.write(struct.pack("<Q", A.size)); f.write(A.tobytes())
        f.write(struct.pack("<Q", 0))
        f.write(struct.pack("<ii?", d, d, True))


# ----------------------------------------------------------------------- Port
class Port:
    """Plain-C restatement. Index data is passed as a dict (see ``index_dict``)."""

    def __init__(self):
        if not os.path.exists(PORT_SO):
            build(port=True, ref=False)
        L = self.L = C.CDLL(PORT_SO)
        L.port_fvec_L2sqr16.restype = C.c_float
        L.port_fvec_L2sqr16.argtypes = [f32p, f32p, sz]
        L.port_fvec_inner_product.restype = C.c_float
        L.port_fvec_inner_product.argtypes = [f32p, f32p, sz]
        L.port_fvec_norm_L2sqr.restype = C.c_float
        L.port_fvec_norm_L2sqr.argtypes = [f32p, sz]
        L.port_opq_apply.argtypes = [sz, sz, f32p, f32p, f32p]
        L.port_inner_prod_table.argtypes = [f32p, f32p, sz, sz, sz, f32p]
        L.port_coarse_exact.argtypes = [f32p, f32p, sz, sz, sz, f32p, u32p]
        L.port_assign_exact.argtypes = [sz, f32p, f32p, sz, sz, u32p]
        L.port_search_adc.argtypes = [sz, sz, sz, sz, sz, sz, f32p, f32p, u32p, f32p, f32p, f32p, i64p, u32p,
                                      u8p, u8p, f32p, C.c_void_p, C.c_void_p]
        L.port_search_grouping.argtypes = [sz, sz, sz, sz, sz, sz, C.c_int, sz, f32p, f32p, u32p, f32p, f32p,
                                           f32p, f32p, i64p, u32p, u8p, u8p, u32p, u32p, f32p, f32p, f32p,
                                           f32p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.port_pq_encode.argtypes = [sz, sz, sz, sz, f32p, f32p, u8p]
        L.port_add_batch_encode.argtypes = [sz, sz, sz, sz, f32p, f32p, C.c_void_p, f32p, f32p, u32p, u8p, u8p,
                                            f32p]
        L.port_compute_alpha.restype = C.c_float
        L.port_compute_alpha.argtypes = [sz, sz, f32p, f32p, f32p, f32p, sz]
        L.port_add_group.argtypes = [sz, sz, sz, sz, f32p, f32p, C.c_void_p, f32p, C.c_uint32, u32p, f32p, sz,
                                     f32p, u32p, f32p, u32p, u32p, u8p, u8p, u32p]
        L.port_centroid_neighbours.argtypes = [f32p, sz, sz, sz, C.c_uint32, u32p, f32p]
        L.port_centroid_norms.argtypes = [f32p, sz, sz, f32p]
        L.port_inter_centroid_dists.argtypes = [f32p, sz, sz, sz, u32p, f32p]

    # -- small kernels
    def l2sqr16(self, x, y):
        return float(self.L.port_fvec_L2sqr16(_f(x), _f(y), len(x)))

    def opq_apply(self, A, x):
        x = _f(x).reshape(-1, A.shape[0])
        out = np.empty_like(x)
        self.L.port_opq_apply(x.shape[0], A.shape[0], _f(A), x, out)
        return out

    def lut(self, q, pqc, M):
        d = q.size
        out = np.empty(M * 256, np.float32)
        self.L.port_inner_prod_table(_f(q), _f(pqc), M, 256, d // M, out)
        return out

    def coarse(self, Q, centroids, nprobe):
        Q = _f(Q).reshape(-1, centroids.shape[1])
        nq = Q.shape[0]
        dists = np.zeros((nq, nprobe), np.float32)
        idxs = np.zeros((nq, nprobe), np.uint32)
        cen = _f(centroids)
        for i in range(nq):
            self.L.port_coarse_exact(Q[i], cen, cen.shape[0], cen.shape[1], nprobe, dists[i], idxs[i])
        return dists, idxs

    def assign(self, X, centroids):
        X = _f(X)
        cen = _f(centroids)
        out = np.zeros(X.shape[0], np.uint32)
        self.L.port_assign_exact(X.shape[0], X, cen, cen.shape[0], cen.shape[1], out)
        return out

    # -- search
    def search(self, ix, Q, coarse_dists, coarse_idxs, k, nprobe, max_codes, do_pruning=False, debug=False):
        """ix: index dict. Q must be the *unrotated* queries; OPQ is applied here
        (IndexIVF_HNSW.cpp:240) when ix['opq'] is not None. Returns heap-ordered
        (distances, labels[, thresholds, scanned, ncode])."""
        d, M = ix["d"], ix["M"]
        Q = _f(Q).reshape(-1, d)
        nq = Q.shape[0]
        if ix.get("opq") is not None:
            Q = self.opq_apply(ix["opq"], Q)
        D = np.zeros((nq, k), np.float32)
        Lb = np.zeros((nq, k), np.int64)
        cd = _f(coarse_dists).reshape(nq, -1)
        ci = np.ascontiguousarray(coarse_idxs, np.uint32).reshape(nq, -1)
        grouping = ix.get("nsubc", 0) > 0
        thr = np.zeros(nq, np.float32)
        nsubc = ix.get("nsubc", 0)
        scanned = np.zeros((nq, nprobe, max(nsubc, 1)), np.uint8)
        ncode = np.zeros(nq, np.uint64)
        scratch = np.zeros(ix["nc"], np.float32)
        for i in range(nq):
            if grouping:
                self.L.port_search_grouping(
                    d, M, 256, nsubc, nprobe, max_codes, int(do_pruning), k, Q[i], cd[i], ci[i],
                    ix["centroids"], ix["pqc"], ix["norm_table"], ix["centroid_norms"], ix["list_off"],
                    ix["ids"], ix["codes"], ix["norm_codes"], ix["nn_idx"], ix["sub_sizes"], ix["alphas"],
                    ix["icd"], scratch, D[i], Lb[i].ctypes.data, thr[i:].ctypes.data, scanned[i].ctypes.data,
                    ncode[i:].ctypes.data)
            else:
                self.L.port_search_adc(
                    d, M, 256, nprobe, max_codes, k, Q[i], cd[i], ci[i], ix["pqc"], ix["norm_table"],
                    ix["centroid_norms"], ix["list_off"], ix["ids"], ix["codes"], ix["norm_codes"], D[i],
                    Lb[i].ctypes.data, ncode[i:].ctypes.data)
        if debug:
            return D, Lb, thr, scanned, ncode
        return D, Lb

    # -- indexing
    def add_batch_encode(self, ix, X, idx):
        X = _f(X)
        n = X.shape[0]
        codes = np.zeros((n, ix["M"]), np.uint8)
        ncodes = np.zeros(n, np.uint8)
        rn = np.zeros(n, np.float32)
        A = ix.get("opq")
        self.L.port_add_batch_encode(n, ix["d"], ix["M"], 256, ix["pqc"], ix["norm_table"],
                                     None if A is None else _f(A).ctypes.data, ix["centroids"], X,
                                     np.ascontiguousarray(idx, np.uint32), codes, ncodes, rn)
        return codes, ncodes, rn

    def centroid_neighbours(self, centroids, nsubc, g):
        cen = _f(centroids)
        nn = np.zeros(nsubc, np.uint32)
        cvn = np.zeros(nsubc, np.float32)
        self.L.port_centroid_neighbours(cen, cen.shape[0], cen.shape[1], nsubc, g, nn, cvn)
        return nn, cvn

    def add_group(self, ix, g, nn, cvn, X, ids):
        X = _f(X).reshape(-1, ix["d"])
        gs = X.shape[0]
        nsubc, M = ix["nsubc"], ix["M"]
        alpha = np.zeros(1, np.float32)
        sub_idx = np.zeros(max(gs, 1), np.uint32)
        oid = np.zeros(max(gs, 1), np.uint32)
        oc = np.zeros((max(gs, 1), M), np.uint8)
        on = np.zeros(max(gs, 1), np.uint8)
        ss = np.zeros(nsubc, np.uint32)
        A = ix.get("opq")
        self.L.port_add_group(ix["d"], M, 256, nsubc, ix["pqc"], ix["norm_table"],
                              None if A is None else _f(A).ctypes.data, ix["centroids"], g,
                              np.ascontiguousarray(nn, np.uint32), _f(cvn), gs, X if gs else np.zeros((1, ix["d"]), np.float32),
                              np.ascontiguousarray(ids, np.uint32) if gs else np.zeros(1, np.uint32),
                              alpha, sub_idx, oid, oc, on, ss)
        return float(alpha[0]), sub_idx[:gs], oid[:gs], oc[:gs], on[:gs], ss

    def centroid_norms(self, centroids):
        cen = _f(centroids)
        out = np.zeros(cen.shape[0], np.float32)
        self.L.port_centroid_norms(cen, cen.shape[0], cen.shape[1], out)
        return out

    def inter_centroid_dists(self, centroids, nn_idx):
        cen = _f(centroids)
        nn = np.ascontiguousarray(nn_idx, np.uint32)
        out = np.zeros(nn.shape, np.float32)
        self.L.port_inter_centroid_dists(cen, cen.shape[0], cen.shape[1], nn.shape[1], nn, out)
        return out


def _f(a):
    return np.ascontiguousarray(a, dtype=np.float32)


# ------------------------------------------------------------------------ Ref
class Ref:
    """The compiled reference (IndexIVF_HNSW / IndexIVF_HNSW_Grouping) via oracle/ref_shim.cpp."""

    def __init__(self, d, nc, code_size, nsubc=0, shipped=False):
        so = REF_SHIPPED_SO if shipped else REF_SO
        if not os.path.exists(so):
            raise RuntimeError("reference library not built: run `make -C oracle ref` where /root/reference exists")
        L = self.L = C.CDLL(so)
        self.d, self.nc, self.M, self.nsubc = d, nc, code_size, nsubc
        L.ref_create.restype = C.c_void_p
        L.ref_create.argtypes = [sz, sz, sz, sz]
        L.ref_destroy.argtypes = [C.c_void_p]
        L.ref_build_quantizer.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p, sz, sz]
        L.ref_set_params.argtypes = [C.c_void_p, sz, sz, sz, C.c_int]
        L.ref_set_do_opq.argtypes = [C.c_void_p, C.c_int]
        L.ref_load_pq.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p]
        L.ref_save_pq.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p]
        L.ref_train_pq.argtypes = [C.c_void_p, sz, f32p]
        L.ref_assign.argtypes = [C.c_void_p, sz, f32p, u32p]
        L.ref_coarse.restype = sz
        L.ref_coarse.argtypes = [C.c_void_p, f32p, sz, f32p, u32p]
        L.ref_add_batch.argtypes = [C.c_void_p, sz, f32p, u32p, u32p]
        L.ref_add_group.argtypes = [C.c_void_p, sz, sz, f32p, u32p]
        for fn in ("ref_compute_centroid_norms", "ref_compute_inter_centroid_dists", "ref_rotate_quantizer"):
            getattr(L, fn).argtypes = [C.c_void_p]
        L.ref_write.argtypes = [C.c_void_p, C.c_char_p]
        L.ref_read.argtypes = [C.c_void_p, C.c_char_p]
        L.ref_search.argtypes = [C.c_void_p, sz, sz, f32p, f32p, i64p, C.c_void_p, C.c_void_p]
        L.ref_search2.argtypes = [C.c_void_p, sz, sz, f32p, f32p, i64p, f32p, u32p]
        L.ref_list_size.restype = sz
        L.ref_list_size.argtypes = [C.c_void_p, sz]
        L.ref_get_list.argtypes = [C.c_void_p, sz, u32p, u8p, u8p]
        L.ref_get_centroid.argtypes = [C.c_void_p, sz, f32p]
        L.ref_get_centroid_norms.argtypes = [C.c_void_p, f32p]
        L.ref_get_pq.argtypes = [C.c_void_p, f32p, f32p]
        L.ref_get_opq.restype = C.c_int
        L.ref_get_opq.argtypes = [C.c_void_p, f32p]
        L.ref_get_group_meta.restype = C.c_int
        L.ref_get_group_meta.argtypes = [C.c_void_p, sz, u32p, u32p, f32p, f32p]
        L.ref_fvec_L2sqr.restype = C.c_float
        L.ref_fvec_L2sqr.argtypes = [f32p, f32p, sz]
        self.h = L.ref_create(d, nc, code_size, nsubc)
        self.tmp = tempfile.mkdtemp(prefix="ivfref_")
        self.do_opq = False

    def close(self):
        if self.h:
            self.L.ref_destroy(self.h)
            self.h = None

    def build_quantizer(self, centroids, M=16, efConstruction=500):
        p = os.path.join(self.tmp, "centroids.fvecs")
        write_fvecs(p, centroids)
        self.L.ref_build_quantizer(self.h, p.encode(), os.path.join(self.tmp, "hnsw.info").encode(),
                                   os.path.join(self.tmp, "hnsw.edges").encode(), M, efConstruction)

    def set_params(self, nprobe, max_codes, efSearch, do_pruning=False):
        self.nprobe = nprobe
        self.L.ref_set_params(self.h, nprobe, max_codes, efSearch, int(do_pruning))

    def set_do_opq(self, on):
        self.do_opq = bool(on)
        self.L.ref_set_do_opq(self.h, int(on))

    def train_pq(self, X):
        X = _f(X)
        self.L.ref_train_pq(self.h, X.shape[0], X)

    def load_pq(self, pqc, norm_table, A=None):
        p1, p2, p3 = (os.path.join(self.tmp, n) for n in ("pq.pq", "norm.pq", "opq.mat"))
        write_pq(p1, self.d, self.M, pqc)
        write_pq(p2, 1, 1, norm_table)
        if A is not None:
            write_opq(p3, A)
        self.L.ref_load_pq(self.h, p1.encode(), p2.encode(), p3.encode() if A is not None else b"")

    def assign(self, X):
        X = _f(X)
        out = np.zeros(X.shape[0], np.uint32)
        self.L.ref_assign(self.h, X.shape[0], X, out)
        return out

    def coarse(self, Q, nprobe):
        Q = _f(Q).reshape(-1, self.d)
        dists = np.zeros((Q.shape[0], nprobe), np.float32)
        idxs = np.zeros((Q.shape[0], nprobe), np.uint32)
        for i in range(Q.shape[0]):
            got = self.L.ref_coarse(self.h, Q[i], nprobe, dists[i], idxs[i])
            assert got == nprobe, "HNSW returned fewer than nprobe results (efSearch < nprobe?)"
        return dists, idxs

    def add_batch(self, X, ids, idx):
        X = _f(X)
        self.L.ref_add_batch(self.h, X.shape[0], X, np.ascontiguousarray(ids, np.uint32),
                             np.ascontiguousarray(idx, np.uint32))

    def add_group(self, g, X, ids):
        X = _f(X).reshape(-1, self.d)
        n = X.shape[0]
        if n == 0:
            X = np.zeros((1, self.d), np.float32)
            ids = np.zeros(1, np.uint32)
        self.L.ref_add_group(self.h, g, n, X, np.ascontiguousarray(ids, np.uint32))

    def compute_centroid_norms(self):
        self.L.ref_compute_centroid_norms(self.h)

    def compute_inter_centroid_dists(self):
        self.L.ref_compute_inter_centroid_dists(self.h)

    def rotate_quantizer(self):
        self.L.ref_rotate_quantizer(self.h)

    def write(self, path):
        self.L.ref_write(self.h, path.encode())

    def read(self, path):
        self.L.ref_read(self.h, path.encode())

    def search(self, Q, k, want_coarse=True):
        Q = _f(Q).reshape(-1, self.d)
        nq = Q.shape[0]
        D = np.zeros((nq, k), np.float32)
        Lb = np.zeros((nq, k), np.int64)
        cd = np.zeros((nq, self.nprobe), np.float32)
        ci = np.zeros((nq, self.nprobe), np.uint32)
        self.L.ref_search(self.h, nq, k, Q, D, Lb, cd.ctypes.data, ci.ctypes.data)
        return (D, Lb, cd, ci) if want_coarse else (D, Lb)

    def search2(self, Q, k, coarse_dists, coarse_idxs):
        Q = _f(Q).reshape(-1, self.d)
        nq = Q.shape[0]
        D = np.zeros((nq, k), np.float32)
        Lb = np.zeros((nq, k), np.int64)
        self.L.ref_search2(self.h, nq, k, Q, D, Lb, _f(coarse_dists), np.ascontiguousarray(coarse_idxs, np.uint32))
        return D, Lb

    def centroids(self):
        out = np.zeros((self.nc, self.d), np.float32)
        for c in range(self.nc):
            self.L.ref_get_centroid(self.h, c, out[c])
        return out

    def export(self):
        """Flatten the reference's per-list vectors into the index dict used by Port and by the CUDA library."""
        nc, M = self.nc, self.M
        sizes = np.array([self.L.ref_list_size(self.h, c) for c in range(nc)], np.int64)
        off = np.zeros(nc + 1, np.int64)
        np.cumsum(sizes, out=off[1:])
        N = int(off[-1])
        ids = np.zeros(max(N, 1), np.uint32)
        codes = np.zeros((max(N, 1), M), np.uint8)
        ncodes = np.zeros(max(N, 1), np.uint8)
        for c in range(nc):
            if sizes[c]:
                a, b = off[c], off[c + 1]
                self.L.ref_get_list(self.h, c, ids[a:b], codes[a:b], ncodes[a:b])
        pqc = np.zeros(self.d * 256, np.float32)
        nt = np.zeros(256, np.float32)
        self.L.ref_get_pq(self.h, pqc, nt)
        cn = np.zeros(nc, np.float32)
        self.L.ref_get_centroid_norms(self.h, cn)
        A = np.zeros((self.d, self.d), np.float32)
        has = self.L.ref_get_opq(self.h, A)
        ix = dict(d=self.d, nc=nc, M=M, nsubc=self.nsubc, centroids=self.centroids(), pqc=pqc, norm_table=nt,
                  centroid_norms=cn, list_off=off, ids=ids[:N], codes=codes[:N], norm_codes=ncodes[:N],
                  opq=A if (has and self.do_opq) else None)
        if self.nsubc:
            ns = self.nsubc
            nn = np.zeros((nc, ns), np.uint32)
            ss = np.zeros((nc, ns), np.uint32)
            al = np.zeros(nc, np.float32)
            icd = np.zeros((nc, ns), np.float32)
            a1 = np.zeros(1, np.float32)
            for c in range(nc):
                self.L.ref_get_group_meta(self.h, c, nn[c], ss[c], a1, icd[c])
                al[c] = a1[0]
            ix.update(nn_idx=nn, sub_sizes=ss, alphas=al, icd=icd)
        return ix
