// <ivf-hnsw/IndexIVF_HNSW.h> of the B200 drop-in.
//
// Same struct name, namespace, public data members and method signatures as the reference
// (reference IndexIVF_HNSW.h:46-191) so that tests/*.cpp and examples/ compile against it unchanged; every
// method forwards to the C ABI of libivfhnsw_b200.so (include/ivfhnsw_b200.h), i.e. to hand-written sm_100a
// kernels.  There is no CPU implementation behind it: without a B200 the constructor reports the C-ABI
// error and exits like the reference's own fatal paths (exit(1), reference utils.h:91-94).
//
// What differs from the reference, by design:
//   * quantizer is a stub (see hnswlib/hnswalg.h): coarse search is an exact GEMM, efSearch is ignored;
//   * the inverted lists live in HBM; `ids`, `codes`, `norm_codes` stay empty until sync_host_lists();
//   * pq / norm_pq / opq_matrix may be swapped by the caller at any time (the tests do,
//     tests/test_ivfhnsw_sift1b.cpp:54-65); the device copies are re-synced lazily before each use;
//   * search() keeps the one-query signature; search_batch() is the entry point that fills the GPU.
#pragma once
#include <cstdint>
#include <cstdio>
#include <fstream>
#include <iostream>
#include <vector>

#include <faiss/FaissAssert.h>
#include <faiss/Heap.h>
#include <faiss/ProductQuantizer.h>
#include <faiss/VectorTransform.h>
#include <faiss/index_io.h>
#include <faiss/utils.h>
#include <hnswlib/hnswalg.h>

#include "orcv.h"
#include "utils.h"

struct ivfhnsw_b200;   // opaque C-ABI handle (include/ivfhnsw_b200.h)

namespace ivfhnsw {

struct IndexIVF_HNSW {
    typedef uint32_t idx_t;

    size_t d;           // vector dimension
    size_t nc;          // number of centroids
    size_t code_size;   // bytes per code

    hnswlib::HierarchicalNSW *quantizer;
    faiss::ProductQuantizer *pq;        // residual codebook (caller may replace it)
    faiss::ProductQuantizer *norm_pq;   // norm codebook
    faiss::LinearTransform *opq_matrix; // OPQ rotation, may stay null
    bool do_opq;

    size_t nprobe;
    size_t max_codes;

    std::vector<std::vector<idx_t> > ids;          // host mirrors, filled by sync_host_lists()
    std::vector<std::vector<uint8_t> > codes;
    std::vector<std::vector<uint8_t> > norm_codes;

    // information associated with the hdr.vec file of write2 (reference IndexIVF_HNSW.h:68-69)
    orcvhdr_t hdr_idx;

    std::vector<float> &get_centroid_norms();

    // probes of the last search() call, FARTHEST first like the reference's TRACE_CENTROIDS vectors
    // (IndexIVF_HNSW.h:73-77, IndexIVF_HNSW.cpp:249-256); filled on demand by trace_centroids()
    std::vector<float> trace_query_centroid_dists;
    std::vector<idx_t> trace_centroid_idxs;
    void trace_centroids(size_t idx_q, bool missed);

public:
    explicit IndexIVF_HNSW(size_t dim, size_t ncentroids, size_t bytes_per_code, size_t nbits_per_idx,
                           size_t max_group_size = 65536);
    virtual ~IndexIVF_HNSW();

    void build_quantizer(const char *path_data, const char *path_info, const char *path_edges, size_t M = 16,
                         size_t efConstruction = 500);
    void assign(size_t n, const float *x, idx_t *labels, size_t k = 1);
    virtual void search(size_t k, const float *x, float *distances, long *labels);
    virtual void add_batch(size_t n, const float *x, const idx_t *xids, const idx_t *precomputed_idx = nullptr);
    virtual void train_pq(size_t n, const float *x);
    virtual void write(const char *path_index);
    virtual void write(const char *path_index, bool do_trunc);
    virtual void read(const char *path);
    void compute_centroid_norms();
    void rotate_quantizer();
    void search2(size_t k, const float *x, float *distances, long *labels, float *query_centroid_dists,
                 idx_t *centroid_idxs);

    // ---- fork extras so that the orca / disk / vector_add / mkindex mains compile and link (thin host wrappers over the
    // same C ABI; reference IndexIVF_HNSW.h:151,167)
    // add_batch + the batch's records [uint64 eid][code_size bytes][norm code] packed into obuf (IndexIVF_HNSW.cpp:140-204)
    virtual void add_batch2(size_t n, const float *x, const idx_t *xids, const idx_t *idx, uint64_t *eids, char *obuf);
    // ORCV directory: hdr.vec, cnorms.vec, pq.vec, normpq.vec, (opq.vec), edge.vec (IndexIVF_HNSW.cpp:672-755)
    virtual void write2(const char *home_dir, size_t n_vecs, bool do_opq, const char *path_edge);
    virtual void search_debug(size_t k, const float *x, float *distances, long *labels) { search(k, x, distances, labels); }
    virtual idx_t search_enn(const float *x, float *distances, long *labels);

    // ---- B200 extensions
    // nq queries at once; rows of `distances`/`labels` come back in faiss max-heap order like search()
    // (sorted = true: ascending instead)
    void search_batch(size_t nq, size_t k, const float *x, float *distances, long *labels, bool sorted = false);
    void assign_u8(size_t n, const uint8_t *x, idx_t *labels);   // bvecs rows without the host float expansion
    void sync_host_lists();                                      // pull ids/codes/norm_codes back from HBM
    ivfhnsw_b200 *handle() { return h_; }

protected:
    int copy_file(const char *file_src, const char *file_dst);
    size_t M;                 // HNSW edges per point as passed to build_quantizer
    float dmatch = 4444.0;    // reference IndexIVF_HNSW.h:90-91
    float dnear = 8888.0;
    std::vector<float> last_query_;   // copy of the last search() query (for trace_centroids)
    std::vector<float> centroid_norms;
    ivfhnsw_b200 *h_;
    size_t nsubc_;
    // fingerprints of the caller-visible quantizers as last uploaded
    uint64_t fp_pq_, fp_norm_, fp_opq_;
    size_t sent_nprobe_, sent_max_codes_;
    int sent_flags_;
    bool pruning_flag_;

    IndexIVF_HNSW(size_t dim, size_t ncentroids, size_t bytes_per_code, size_t nbits_per_idx, size_t nsubc, int);
    void init(size_t dim, size_t ncentroids, size_t bytes_per_code, size_t nbits_per_idx, size_t nsubc);
    void sync_quantizers();            // re-upload pq / norm_pq / opq_matrix / parameters if they changed
    void pull_quantizers();            // after train_pq: device -> pq / norm_pq / opq_matrix
    void ck(int rc, const char *what) const;   // C-ABI error -> message + exit(1)
};

}  // namespace ivfhnsw
