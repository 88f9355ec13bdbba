"""ivf-hnsw_b200 -- host-side Python mirror of the reference's index API over the
C ABI of libivfhnsw_b200.so (include/ivfhnsw_b200.h).

The product is the CUDA library; this module only binds it (ctypes) with the
reference's names and argument meaning:

    IndexIVF_HNSW(dim, ncentroids, bytes_per_code, nbits_per_idx)         IndexIVF_HNSW.h:94-95
    IndexIVF_HNSW_Grouping(dim, ncentroids, bytes_per_code, nbits, nsubc)  IndexIVF_HNSW_Grouping.h:25-26
    .build_quantizer / .assign / .train_pq* / .add_batch / .add_group /
    .compute_centroid_norms / .compute_inter_centroid_dists / .rotate_quantizer /
    .search / .search2 / .read / .write          (IndexIVF_HNSW.h:107-179)
    public members nprobe, max_codes, do_opq, do_pruning

There is no CPU fallback: importing works anywhere (so that the ABI can be
checked), every compute call needs an sm_100 GPU and raises otherwise.
"""
import ctypes as C
import os
import struct

import numpy as np

__all__ = ["IndexIVF_HNSW", "IndexIVF_HNSW_Grouping", "lib", "LIB_PATH", "Stats", "B200Error", "build_library"]

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libivfhnsw_b200.so")


class B200Error(RuntimeError):
    pass


def build_library(force=False):
    """Compile csrc/ for sm_100a in-tree (nvcc cross-compiles without a GPU)."""
    import subprocess
    if force:
        subprocess.check_call(["make", "-s", "-C", os.path.join(HERE, "csrc"), "clean"])
    subprocess.check_call(["make", "-s", "-C", os.path.join(HERE, "csrc"), "-j8"])


class Stats(C.Structure):
    _fields_ = [("kernel_launches", C.c_uint64), ("scanned_codes", C.c_uint64), ("walked_codes", C.c_uint64),
                ("queries", C.c_uint64), ("scan_ms", C.c_double), ("coarse_ms", C.c_double),
                ("prep_ms", C.c_double), ("scan_launches", C.c_uint64), ("coarse_flops", C.c_uint64),
                ("gemm_ms", C.c_double), ("g1_ms", C.c_double), ("lut_ms", C.c_double), ("g1_rows", C.c_uint64),
                ("prep_launches", C.c_uint64)]


_lib = None


def lib():
    """The loaded C ABI. Fails loudly when the extension has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B200Error("libivfhnsw_b200.so is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                        "(there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    H, sz, vp = C.c_void_p, C.c_size_t, C.c_void_p
    sigs = {
        "ivfhnsw_b200_abi_version": ([], C.c_int),
        "ivfhnsw_b200_create": ([C.POINTER(H), sz, sz, sz, sz, sz, C.c_int], C.c_int),
        "ivfhnsw_b200_destroy": ([H], None),
        "ivfhnsw_b200_last_error": ([H], C.c_char_p),
        "ivfhnsw_b200_set_shard": ([H, sz, sz], C.c_int),
        "ivfhnsw_b200_set_centroids": ([H, vp], C.c_int),
        "ivfhnsw_b200_set_pq": ([H, vp], C.c_int),
        "ivfhnsw_b200_set_norm_pq": ([H, vp], C.c_int),
        "ivfhnsw_b200_set_opq": ([H, vp], C.c_int),
        "ivfhnsw_b200_set_params": ([H, sz, sz, C.c_int, C.c_int], C.c_int),
        "ivfhnsw_b200_compute_centroid_norms": ([H], C.c_int),
        "ivfhnsw_b200_get_centroid_norms": ([H, vp], C.c_int),
        "ivfhnsw_b200_set_centroid_norms": ([H, vp], C.c_int),
        "ivfhnsw_b200_rotate_quantizer": ([H], C.c_int),
        "ivfhnsw_b200_get_centroids": ([H, vp], C.c_int),
        "ivfhnsw_b200_train_pq": ([H, sz, vp], C.c_int),
        "ivfhnsw_b200_get_pq": ([H, vp, vp, vp], C.c_int),
        "ivfhnsw_b200_train_centroids": ([H, sz, vp, C.c_int, C.c_long, C.POINTER(C.c_double), C.POINTER(sz)], C.c_int),
        "ivfhnsw_b200_assign": ([H, sz, vp, vp, sz], C.c_int),
        "ivfhnsw_b200_assign_u8": ([H, sz, vp, vp], C.c_int),
        "ivfhnsw_b200_assign_dev": ([H, sz, vp, vp], C.c_int),
        "ivfhnsw_b200_add_batch": ([H, sz, vp, vp, vp], C.c_int),
        "ivfhnsw_b200_add_batch_codes": ([H, sz, vp, vp, vp, vp, vp], C.c_int),
        "ivfhnsw_b200_rerank": ([H, sz, sz, sz, vp, vp, vp, vp, vp], C.c_int),
        "ivfhnsw_b200_add_group": ([H, sz, sz, vp, vp], C.c_int),
        "ivfhnsw_b200_add_points_dev": ([H, sz, vp, vp, vp], C.c_int),
        "ivfhnsw_b200_finalize": ([H], C.c_int),
        "ivfhnsw_b200_flush_groups": ([H], C.c_int),
        "ivfhnsw_b200_load_lists": ([H, vp, vp, vp, vp, vp, vp, vp, vp], C.c_int),
        "ivfhnsw_b200_ntotal": ([H, C.POINTER(sz)], C.c_int),
        "ivfhnsw_b200_list_sizes": ([H, vp], C.c_int),
        "ivfhnsw_b200_get_list_meta_dev": ([H, vp, vp, vp], C.c_int),
        "ivfhnsw_b200_set_list_meta_dev": ([H, vp, vp, vp], C.c_int),
        "ivfhnsw_b200_export_lists": ([H, vp, vp, vp, vp, vp, vp, vp, vp], C.c_int),
        "ivfhnsw_b200_write_index": ([H, C.c_char_p], C.c_int),
        "ivfhnsw_b200_read_index": ([H, C.c_char_p], C.c_int),
        "ivfhnsw_b200_coarse": ([H, sz, vp, sz, vp, vp], C.c_int),
        "ivfhnsw_b200_search": ([H, sz, sz, vp, vp, vp], C.c_int),
        "ivfhnsw_b200_search_given_coarse": ([H, sz, sz, vp, vp, vp, vp, vp], C.c_int),
        "ivfhnsw_b200_search_dev": ([H, sz, sz, vp, vp, vp], C.c_int),
        "ivfhnsw_b200_search_given_coarse_dev": ([H, sz, sz, vp, vp, vp, vp, vp], C.c_int),
        "ivfhnsw_b200_coarse_dev": ([H, sz, vp, sz, vp, vp], C.c_int),
        "ivfhnsw_b200_coarse_noalloc_dev": ([H, sz, vp, sz, vp, vp, vp], C.c_int),
        "ivfhnsw_b200_search_partial_dev": ([H, sz, sz, vp, vp, vp, vp, vp, vp], C.c_int),
        "ivfhnsw_b200_merge_partials_dev": ([H, sz, sz, sz, vp, vp, vp, vp, vp], C.c_int),
        "ivfhnsw_b200_tables_bytes": ([H, C.POINTER(C.c_size_t)], C.c_int),
        "ivfhnsw_b200_prepare_dev": ([H, sz, vp, sz, sz, vp], C.c_int),
        "ivfhnsw_b200_scan_tables_dev": ([H, sz, sz, vp, vp, vp, vp, vp], C.c_int),
        "ivfhnsw_b200_pack_records_dev": ([H, sz, vp, sz, vp, vp, vp, vp], C.c_int),
        "ivfhnsw_b200_scan_packed_dev": ([H, sz, sz, vp, vp, vp, vp, vp, vp, vp], C.c_int),
        "ivfhnsw_b200_sync": ([H], C.c_int),
        "ivfhnsw_b200_set_stream": ([H, vp], C.c_int),
        "ivfhnsw_b200_set_timing": ([H, C.c_int], C.c_int),
        "ivfhnsw_b200_get_stats": ([H, C.POINTER(Stats)], C.c_int),
        "ivfhnsw_b200_reset_stats": ([H], C.c_int),
        "ivfhnsw_b200_last_thresholds": ([H, sz, vp, vp], C.c_int),
        "ivfhnsw_b200_set_option": ([H, C.c_char_p, C.c_long], C.c_int),
    }
    for name, (args, res) in sigs.items():
        fn = getattr(L, name)  # AttributeError here == a symbol the header declares is not exported
        fn.argtypes = args
        fn.restype = res
    _lib = L
    return L


def _np(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    return int(a.data_ptr())  # torch tensor (device or pinned host)


# ------------------------------------------------------------------ faiss-compatible file formats
def read_ProductQuantizer(path):
    """faiss::read_ProductQuantizer (faiss/index_io.cpp:504-513): size_t d, M, nbits; vector<float> centroids."""
    with open(path, "rb") as f:
        d, M, nbits, n = struct.unpack("<QQQQ", f.read(32))
        c = np.frombuffer(f.read(4 * n), dtype=np.float32).copy()
    return d, M, nbits, c


def write_ProductQuantizer(path, d, M, centroids, nbits=8):
    c = _np(centroids, np.float32).ravel()
    with open(path, "wb") as f:
        f.write(struct.pack("<QQQQ", d, M, nbits, c.size))
        f.write(c.tobytes())


def read_VectorTransform(path):
    """faiss::read_VectorTransform for a LinearTransform ('LTra', faiss/index_io.cpp:425-470) -> A [d_out][d_in]."""
    with open(path, "rb") as f:
        if f.read(4) != b"LTra":
            raise B200Error("only generic LinearTransform (OPQ) files are supported")
        (have_bias,) = struct.unpack("<?", f.read(1))
        (na,) = struct.unpack("<Q", f.read(8))
        A = np.frombuffer(f.read(4 * na), dtype=np.float32).copy()
        (nb,) = struct.unpack("<Q", f.read(8))
        f.read(4 * nb)
        d_in, d_out = struct.unpack("<ii", f.read(8))
    if have_bias:
        raise B200Error("LinearTransform with bias is not used by ivf-hnsw")
    return A.reshape(d_out, d_in)


def write_VectorTransform(path, A):
    A = _np(A, np.float32)
    with open(path, "wb") as f:
        f.write(b"LTra")
        f.write(struct.pack("<?", False))
        f.write(struct.pack("<Q", A.size)); f.write(A.tobytes())
        f.write(struct.pack("<Q", 0))
        f.write(struct.pack("<ii?", A.shape[1], A.shape[0], True))


def read_fvecs(path, d, n=None):
    a = np.fromfile(path, dtype=np.float32)
    a = a.reshape(-1, d + 1)
    if n is not None:
        a = a[:n]
    if not np.all(a[:, 0].view(np.uint32) == d):
        raise B200Error("file error")  # utils.h:91-94
    return np.ascontiguousarray(a[:, 1:])


class IndexIVF_HNSW:
    """IVFADC index (IndexIVF_HNSW.h:46-191) running on one B200."""

    def __init__(self, dim, ncentroids, bytes_per_code, nbits_per_idx=8, nsubc=0, device=0):
        self.d, self.nc, self.code_size, self.nsubc = int(dim), int(ncentroids), int(bytes_per_code), int(nsubc)
        self.nprobe, self.max_codes = 1, 10 ** 9
        self.do_opq, self.do_pruning = False, False
        self.pq_centroids = None      # index->pq->centroids
        self.norm_centroids = None    # index->norm_pq->centroids
        self.opq_matrix = None        # index->opq_matrix->A
        self.device, self._ext_stream = int(device), None
        self._L = lib()
        self._h = C.c_void_p()
        rc = self._L.ivfhnsw_b200_create(C.byref(self._h), self.d, self.nc, self.code_size, nbits_per_idx,
                                         self.nsubc, device)
        if rc != 0:
            msg = self._L.ivfhnsw_b200_last_error(self._h).decode() if self._h else "create failed"
            if self._h:
                self._L.ivfhnsw_b200_destroy(self._h)
                self._h = None
            raise B200Error(msg)

    # -- plumbing
    def _ck(self, rc):
        if rc < 0:
            raise B200Error(self._L.ivfhnsw_b200_last_error(self._h).decode())
        return rc

    def close(self):
        if getattr(self, "_h", None):
            self._L.ivfhnsw_b200_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _push_params(self):
        self._ck(self._L.ivfhnsw_b200_set_params(self._h, self.nprobe, self.max_codes, int(self.do_opq),
                                                 int(self.do_pruning)))

    def set_option(self, name, value):
        self._ck(self._L.ivfhnsw_b200_set_option(self._h, name.encode(), int(value)))

    def set_shard(self, c_begin, c_end):
        self._ck(self._L.ivfhnsw_b200_set_shard(self._h, c_begin, c_end))

    # -- model state
    def build_quantizer(self, path_data=None, path_info=None, path_edges=None, M=16, efConstruction=500,
                        centroids=None):
        """IndexIVF_HNSW::build_quantizer (IndexIVF_HNSW.cpp:40-65). The coarse quantizer here is the exact
        GEMM search over the centroid table, so only the centroid file is read; no graph is built."""
        if centroids is None:
            centroids = read_fvecs(path_data, self.d, self.nc)
        c = _np(centroids, np.float32)
        if c.shape != (self.nc, self.d):
            raise B200Error("centroid table has shape %s, expected (%d, %d)" % (c.shape, self.nc, self.d))
        self._ck(self._L.ivfhnsw_b200_set_centroids(self._h, c.ctypes.data))

    def set_pq(self, pq_centroids, norm_centroids, opq_matrix=None):
        """What the tests do by assigning index->pq / norm_pq / opq_matrix (tests/test_ivfhnsw_sift1b.cpp:54-65)."""
        self.pq_centroids = _np(pq_centroids, np.float32).ravel()
        self.norm_centroids = _np(norm_centroids, np.float32).ravel()
        if self.pq_centroids.size != self.d * 256 or self.norm_centroids.size != 256:
            raise B200Error("bad codebook sizes")
        self._ck(self._L.ivfhnsw_b200_set_pq(self._h, self.pq_centroids.ctypes.data))
        self._ck(self._L.ivfhnsw_b200_set_norm_pq(self._h, self.norm_centroids.ctypes.data))
        self.opq_matrix = None if opq_matrix is None else _np(opq_matrix, np.float32)
        self._ck(self._L.ivfhnsw_b200_set_opq(self._h, None if self.opq_matrix is None else
                                              self.opq_matrix.ctypes.data))

    def train_pq(self, x):
        """IndexIVF_HNSW::train_pq / IndexIVF_HNSW_Grouping::train_pq on the GPU; afterwards
        pq_centroids / norm_centroids / opq_matrix hold the trained model like index->pq etc."""
        self._push_params()
        x = _np(x, np.float32).reshape(-1, self.d)
        self._ck(self._L.ivfhnsw_b200_train_pq(self._h, x.shape[0], x.ctypes.data))
        self.pq_centroids = np.zeros(self.d * 256, np.float32)
        self.norm_centroids = np.zeros(256, np.float32)
        A = np.zeros((self.d, self.d), np.float32) if self.do_opq else None
        self._ck(self._L.ivfhnsw_b200_get_pq(self._h, self.pq_centroids.ctypes.data, self.norm_centroids.ctypes.data,
                                             None if A is None else A.ctypes.data))
        self.opq_matrix = A

    def train_centroids(self, x_dev, niter=10, seed=1234):
        """Coarse k-means for the centroid file on the GPU (faiss::Clustering::train with K-A1 as the assignment
        step, void clusters re-seeded like km_update_centroids).  x_dev: float32 CUDA tensor [n][d].  The result
        becomes the index' centroid table; returns (objective, void clusters re-seeded)."""
        obj, ns = C.c_double(), C.c_size_t()
        with self._ordered():
            self._ck(self._L.ivfhnsw_b200_train_centroids(self._h, x_dev.shape[0], _ptr(x_dev), niter, seed,
                                                          C.byref(obj), C.byref(ns)))
        return obj.value, ns.value

    def compute_centroid_norms(self):
        self._ck(self._L.ivfhnsw_b200_finalize(self._h))  # flush point, SURVEY.md 8(b)
        self._ck(self._L.ivfhnsw_b200_compute_centroid_norms(self._h))

    def compute_inter_centroid_dists(self):
        self._ck(self._L.ivfhnsw_b200_finalize(self._h))  # computed inside finalize for Grouping

    def get_centroid_norms(self):
        out = np.zeros(self.nc, np.float32)
        self._ck(self._L.ivfhnsw_b200_get_centroid_norms(self._h, out.ctypes.data))
        return out

    def get_centroids(self):
        out = np.zeros((self.nc, self.d), np.float32)
        self._ck(self._L.ivfhnsw_b200_get_centroids(self._h, out.ctypes.data))
        return out

    def rotate_quantizer(self):
        if not self.do_opq:
            raise B200Error("OPQ encoding is turned off")  # IndexIVF_HNSW.cpp:790-793 aborts
        self._ck(self._L.ivfhnsw_b200_rotate_quantizer(self._h))

    # -- indexing
    def assign(self, x, k=1):
        if isinstance(x, np.ndarray) and x.dtype == np.uint8:
            x = np.ascontiguousarray(x)
            labels = np.zeros(x.shape[0], np.uint32)
            self._ck(self._L.ivfhnsw_b200_assign_u8(self._h, x.shape[0], x.ctypes.data, labels.ctypes.data))
            return labels
        x = _np(x, np.float32).reshape(-1, self.d)
        labels = np.zeros(x.shape[0], np.uint32)
        self._ck(self._L.ivfhnsw_b200_assign(self._h, x.shape[0], x.ctypes.data, labels.ctypes.data, k))
        return labels

    def assign_dev(self, x_dev, labels_dev):
        """IndexIVF_HNSW::assign for device-resident vectors (torch tensors): labels_dev int32 [n]."""
        with self._ordered():
            self._ck(self._L.ivfhnsw_b200_assign_dev(self._h, x_dev.shape[0], _ptr(x_dev), _ptr(labels_dev)))

    def add_batch(self, x, xids, precomputed_idx=None):
        self._push_params()
        x = _np(x, np.float32).reshape(-1, self.d)
        xids = _np(xids, np.uint32)
        idx = None if precomputed_idx is None else _np(precomputed_idx, np.uint32)
        self._ck(self._L.ivfhnsw_b200_add_batch(self._h, x.shape[0], x.ctypes.data, xids.ctypes.data,
                                                None if idx is None else idx.ctypes.data))

    # -- stream ordering for the device-pointer entry points.  The C ABI enqueues on the handle's stream
    # (private unless set_stream() was called).  Tensors usually come from torch's current stream: unless
    # that IS the handle's stream, wait for it before the call and drain the handle's stream after it,
    # so that the *_dev methods are safe by default and fully asynchronous once the stream is shared.
    def _ordered(self):
        idx = self

        class _Ctx:
            def __enter__(self_c):
                self_c.shared = False
                try:
                    import torch
                    cur = torch.cuda.current_stream(idx.device).cuda_stream
                    self_c.shared = idx._ext_stream is not None and idx._ext_stream == cur
                    if not self_c.shared:
                        torch.cuda.current_stream(idx.device).synchronize()
                except ImportError:
                    pass

            def __exit__(self_c, et, ev, tb):
                if et is None and not self_c.shared:
                    idx.sync()
                return False
        return _Ctx()

    def add_points_dev(self, x_dev, ids_dev, idx_dev=None):
        """Bulk ingest of device-resident vectors (torch tensors); idx_dev=None assigns on the GPU."""
        self._push_params()
        with self._ordered():
            self._ck(self._L.ivfhnsw_b200_add_points_dev(self._h, x_dev.shape[0], _ptr(x_dev), _ptr(ids_dev),
                                                         _ptr(idx_dev)))

    def flush_groups(self):
        """Grouping: the groups staged so far are complete -- encode them and drop their raw vectors."""
        self._push_params()
        self._ck(self._L.ivfhnsw_b200_flush_groups(self._h))

    def finalize(self):
        self._push_params()
        self._ck(self._L.ivfhnsw_b200_finalize(self._h))

    @property
    def ntotal(self):
        n = C.c_size_t()
        self._L.ivfhnsw_b200_ntotal(self._h, C.byref(n))
        return n.value

    # -- lists
    def load_index_dict(self, ix):
        """Load the flattened per-list vectors (what read() fills, IndexIVF_HNSW.cpp:758-779)."""
        off = _np(ix["list_off"], np.int64)
        ids = _np(ix["ids"], np.uint32)
        codes = _np(ix["codes"], np.uint8)
        ncodes = _np(ix["norm_codes"], np.uint8)
        g = self.nsubc > 0
        nn = _np(ix["nn_idx"], np.uint32) if g else None
        ss = _np(ix["sub_sizes"], np.uint32) if g else None
        al = _np(ix["alphas"], np.float32) if g else None
        icd = _np(ix["icd"], np.float32) if g else None
        self._keep = (off, ids, codes, ncodes, nn, ss, al, icd)
        self._ck(self._L.ivfhnsw_b200_load_lists(self._h, off.ctypes.data, ids.ctypes.data, codes.ctypes.data,
                                                 ncodes.ctypes.data, _ptr(nn), _ptr(ss), _ptr(al), _ptr(icd)))
        self._keep = None       # load_lists returns after its copies have completed: the host arrays are free again
        if "centroid_norms" in ix:
            cn = _np(ix["centroid_norms"], np.float32)
            self._ck(self._L.ivfhnsw_b200_set_centroid_norms(self._h, cn.ctypes.data))

    def export_index_dict(self):
        self.finalize()
        n, nc, M, ns = self.ntotal, self.nc, self.code_size, self.nsubc
        off = np.zeros(nc + 1, np.int64)
        ids = np.zeros(max(n, 1), np.uint32)
        codes = np.zeros((max(n, 1), M), np.uint8)
        ncodes = np.zeros(max(n, 1), np.uint8)
        nn = np.zeros((nc, max(ns, 1)), np.uint32)
        ss = np.zeros((nc, max(ns, 1)), np.uint32)
        al = np.zeros(nc, np.float32)
        icd = np.zeros((nc, max(ns, 1)), np.float32)
        self._ck(self._L.ivfhnsw_b200_export_lists(self._h, off.ctypes.data, ids.ctypes.data, codes.ctypes.data,
                                                   ncodes.ctypes.data, nn.ctypes.data, ss.ctypes.data,
                                                   al.ctypes.data, icd.ctypes.data))
        out = dict(d=self.d, nc=nc, M=M, nsubc=ns, list_off=off, ids=ids[:n], codes=codes[:n], norm_codes=ncodes[:n],
                   centroid_norms=self.get_centroid_norms())
        if ns:
            out.update(nn_idx=nn, sub_sizes=ss, alphas=al, icd=icd)
        return out

    def write(self, path):
        """Byte-compatible with IndexIVF_HNSW::write / IndexIVF_HNSW_Grouping::write
        (IndexIVF_HNSW.cpp:637-663, IndexIVF_HNSW_Grouping.cpp:397-438); native (ivfhnsw_b200_write_index)."""
        self._push_params()
        self._ck(self._L.ivfhnsw_b200_write_index(self._h, os.fsencode(path)))

    def read(self, path):
        """IndexIVF_HNSW::read / IndexIVF_HNSW_Grouping::read (IndexIVF_HNSW.cpp:758-779, ...Grouping.cpp:445-483)."""
        self._push_params()
        self._ck(self._L.ivfhnsw_b200_read_index(self._h, os.fsencode(path)))

    # -- search
    def coarse(self, x, nprobe=None):
        self._push_params()
        nprobe = nprobe or self.nprobe
        x = _np(x, np.float32).reshape(-1, self.d)
        dists = np.zeros((x.shape[0], nprobe), np.float32)
        idxs = np.zeros((x.shape[0], nprobe), np.uint32)
        self._ck(self._L.ivfhnsw_b200_coarse(self._h, x.shape[0], x.ctypes.data, nprobe, dists.ctypes.data,
                                             idxs.ctypes.data))
        return dists, idxs

    def search(self, k, x):
        """search(k, x, distances, labels) for a batch of queries; returns (distances, labels) sorted ascending."""
        self._push_params()
        x = _np(x, np.float32).reshape(-1, self.d)
        D = np.zeros((x.shape[0], k), np.float32)
        L = np.zeros((x.shape[0], k), np.int64)
        self._ck(self._L.ivfhnsw_b200_search(self._h, x.shape[0], k, x.ctypes.data, D.ctypes.data, L.ctypes.data))
        return D, L

    def search2(self, k, x, query_centroid_dists, centroid_idxs):
        """IndexIVF_HNSW::search2 (IndexIVF_HNSW.cpp:453-492): coarse candidates supplied by the caller."""
        self._push_params()
        x = _np(x, np.float32).reshape(-1, self.d)
        cd = _np(query_centroid_dists, np.float32).reshape(x.shape[0], self.nprobe)
        ci = _np(centroid_idxs, np.uint32).reshape(x.shape[0], self.nprobe)
        D = np.zeros((x.shape[0], k), np.float32)
        L = np.zeros((x.shape[0], k), np.int64)
        self._ck(self._L.ivfhnsw_b200_search_given_coarse(self._h, x.shape[0], k, x.ctypes.data, cd.ctypes.data,
                                                          ci.ctypes.data, D.ctypes.data, L.ctypes.data))
        return D, L

    def search_dev(self, k, x_dev, D_dev, L_dev, coarse=None):
        """Device-resident search, asynchronous on the handle's stream (call sync())."""
        self._push_params()
        nq = x_dev.shape[0]
        if coarse is None:
            with self._ordered():
                self._ck(self._L.ivfhnsw_b200_search_dev(self._h, nq, k, _ptr(x_dev), _ptr(D_dev), _ptr(L_dev)))
        else:
            with self._ordered():
                self._ck(self._L.ivfhnsw_b200_search_given_coarse_dev(self._h, nq, k, _ptr(x_dev), _ptr(coarse[0]),
                                                                      _ptr(coarse[1]), _ptr(D_dev), _ptr(L_dev)))

    def coarse_dev(self, x_dev, nprobe, d_dev, i_dev):
        self._push_params()
        with self._ordered():
            self._ck(self._L.ivfhnsw_b200_coarse_dev(self._h, x_dev.shape[0], _ptr(x_dev), nprobe, _ptr(d_dev),
                                                     _ptr(i_dev)))

    def coarse_noalloc_dev(self, xq_dev, nprobe, d_dev, i_dev, flags_dev=None):
        """ivfhnsw_b200_coarse_noalloc_dev: the exact coarse search with its scratch in the library's static device
        array (allocates nothing; queries already in the centroid table's space; see include/ivfhnsw_b200.h)."""
        with self._ordered():
            self._ck(self._L.ivfhnsw_b200_coarse_noalloc_dev(self._h, xq_dev.shape[0], _ptr(xq_dev), nprobe,
                                                             _ptr(d_dev), _ptr(i_dev), _ptr(flags_dev)))

    def rerank(self, k, x, cand, cand_labels):
        """The re-rank half of searchDisk (IndexIVF_HNSW_Grouping.cpp:365-395): exact fvec_L2sqr distances of the
        candidates' raw vectors cand [nq][ncand][d], k best per query ascending by (distance, label)."""
        x = _np(x, np.float32).reshape(-1, self.d)
        cand = _np(cand, np.float32)
        lab = _np(cand_labels, np.int64)
        nq, ncand = lab.shape
        D = np.zeros((nq, k), np.float32)
        L = np.zeros((nq, k), np.int64)
        self._ck(self._L.ivfhnsw_b200_rerank(self._h, nq, ncand, k, x.ctypes.data, cand.ctypes.data, lab.ctypes.data,
                                             D.ctypes.data, L.ctypes.data))
        return D, L

    def add_batch2(self, x, xids, precomputed_idx):
        """add_batch that also returns (codes, norm_codes) of the batch, what IndexIVF_HNSW::add_batch2 packs into obuf."""
        self._push_params()
        x = _np(x, np.float32).reshape(-1, self.d)
        xids, idx = _np(xids, np.uint32), _np(precomputed_idx, np.uint32)
        codes = np.zeros((x.shape[0], self.code_size), np.uint8)
        ncodes = np.zeros(x.shape[0], np.uint8)
        self._ck(self._L.ivfhnsw_b200_add_batch_codes(self._h, x.shape[0], x.ctypes.data, xids.ctypes.data, idx.ctypes.data,
                                                      codes.ctypes.data, ncodes.ctypes.data))
        return codes, ncodes

    def search_host_buffers(self, k, x_host, D_host, L_host):
        """The C-ABI call with caller-owned HOST buffers (numpy or pinned torch): H2D + search + D2H."""
        self._push_params()
        self._ck(self._L.ivfhnsw_b200_search(self._h, x_host.shape[0], k, _ptr(x_host), _ptr(D_host), _ptr(L_host)))

    def search_partial_dev(self, k, x_dev, part_d, part_seq, part_labels, coarse=None):
        """This shard's sorted partial top-k (dist, scan-order key, label) for all queries."""
        self._push_params()
        cd, ci = (None, None) if coarse is None else coarse
        with self._ordered():
            self._ck(self._L.ivfhnsw_b200_search_partial_dev(self._h, x_dev.shape[0], k, _ptr(x_dev), _ptr(cd),
                                                             _ptr(ci), _ptr(part_d), _ptr(part_seq),
                                                             _ptr(part_labels)))

    def tables_bytes(self):
        """Bytes of one per-query work-table record at the current nprobe (see ivfhnsw_b200_prepare_dev)."""
        self._push_params()
        n = C.c_size_t()
        self._ck(self._L.ivfhnsw_b200_tables_bytes(self._h, C.byref(n)))
        return int(n.value)

    def prepare_dev(self, x_dev, q_begin, q_end, tables_dev):
        """Coarse probes + work tables of queries [q_begin, q_end) into their records of tables_dev (uint8 [nq][bytes])."""
        self._push_params()
        with self._ordered():
            self._ck(self._L.ivfhnsw_b200_prepare_dev(self._h, x_dev.shape[0], _ptr(x_dev), q_begin, q_end,
                                                      _ptr(tables_dev)))

    def scan_tables_dev(self, k, x_dev, tables_dev, part_d, part_seq, part_labels):
        """Scan of this shard's lists for all queries, driven by (gathered) records."""
        self._push_params()
        with self._ordered():
            self._ck(self._L.ivfhnsw_b200_scan_tables_dev(self._h, x_dev.shape[0], k, _ptr(x_dev), _ptr(tables_dev),
                                                          _ptr(part_d), _ptr(part_seq), _ptr(part_labels)))

    def pack_records_dev(self, nq, tables_dev, nparts, packed_dev, cnt_dev, off_dev, part_words_dev):
        """Records of nq home queries -> one packed stream per owning rank (ivfhnsw_b200_pack_records_dev)."""
        self._push_params()
        with self._ordered():
            self._ck(self._L.ivfhnsw_b200_pack_records_dev(self._h, nq, _ptr(tables_dev), nparts, _ptr(packed_dev),
                                                           _ptr(cnt_dev), _ptr(off_dev), _ptr(part_words_dev)))

    def scan_packed_dev(self, k, x_dev, packed_dev, qoff_dev, qnp_dev, part_d, part_seq, part_labels):
        """Scan of this shard's lists driven by packed records (ivfhnsw_b200_scan_packed_dev)."""
        self._push_params()
        with self._ordered():
            self._ck(self._L.ivfhnsw_b200_scan_packed_dev(self._h, x_dev.shape[0], k, _ptr(x_dev), _ptr(packed_dev),
                                                          _ptr(qoff_dev), _ptr(qnp_dev), _ptr(part_d), _ptr(part_seq),
                                                          _ptr(part_labels)))

    def get_list_meta_dev(self, list_size_dev, sub_sizes_dev=None, alphas_dev=None):
        """This handle's per-list tables (sharded construction, ivfhnsw_b200_get_list_meta_dev)."""
        with self._ordered():
            self._ck(self._L.ivfhnsw_b200_get_list_meta_dev(self._h, _ptr(list_size_dev), _ptr(sub_sizes_dev),
                                                            _ptr(alphas_dev)))

    def set_list_meta_dev(self, list_size_dev, sub_sizes_dev=None, alphas_dev=None):
        with self._ordered():
            self._ck(self._L.ivfhnsw_b200_set_list_meta_dev(self._h, _ptr(list_size_dev), _ptr(sub_sizes_dev),
                                                            _ptr(alphas_dev)))

    def merge_partials_dev(self, k, nq, nparts, part_d, part_seq, part_labels, D_dev, L_dev):
        with self._ordered():
            self._ck(self._L.ivfhnsw_b200_merge_partials_dev(self._h, nq, k, nparts, _ptr(part_d), _ptr(part_seq),
                                                             _ptr(part_labels), _ptr(D_dev), _ptr(L_dev)))

    def set_stream(self, cuda_stream_ptr):
        """Run on a caller-owned stream, e.g. torch.cuda.current_stream().cuda_stream (0/None = private)."""
        self._ck(self._L.ivfhnsw_b200_set_stream(self._h, cuda_stream_ptr or None))
        self._ext_stream = cuda_stream_ptr or None

    def sync(self):
        self._ck(self._L.ivfhnsw_b200_sync(self._h))

    # -- introspection
    def set_timing(self, on):
        self._ck(self._L.ivfhnsw_b200_set_timing(self._h, int(on)))

    def stats(self):
        s = Stats()
        self._ck(self._L.ivfhnsw_b200_get_stats(self._h, C.byref(s)))
        return s

    def reset_stats(self):
        self._ck(self._L.ivfhnsw_b200_reset_stats(self._h))

    def last_thresholds(self, cap):
        thr = np.zeros(cap, np.float32)
        sc = np.zeros(cap, np.uint32)
        n = self._ck(self._L.ivfhnsw_b200_last_thresholds(self._h, cap, thr.ctypes.data, sc.ctypes.data))
        return thr[:n], sc[:n]


class IndexIVF_HNSW_Grouping(IndexIVF_HNSW):
    """IVFADC + Grouping (+ Pruning) index (IndexIVF_HNSW_Grouping.h:15-75)."""

    def __init__(self, dim, ncentroids, bytes_per_code, nbits_per_idx, nsubcentroids, device=0):
        super().__init__(dim, ncentroids, bytes_per_code, nbits_per_idx, nsubc=nsubcentroids, device=device)

    def add_group(self, group_idx, x, ids):
        self._push_params()
        x = _np(x, np.float32).reshape(-1, self.d)
        ids = _np(ids, np.uint32)
        self._ck(self._L.ivfhnsw_b200_add_group(self._h, group_idx, x.shape[0],
                                                x.ctypes.data if x.shape[0] else None,
                                                ids.ctypes.data if x.shape[0] else None))
