// Host-callable launchers of the ivfhnsw_b200 kernels.  Every launcher enqueues on
// `st`, returns the cudaError of the launch and bumps *launches by the kernels it issued.
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cstddef>
#include <cstdint>

namespace ivfb200 {

// ----------------------------------------------------------------------------- index (device CSR)
struct DeviceLists {
    // slot space: list c occupies slots [list_slot[c], list_slot[c] + list_size[c]); list starts are
    // aligned to `align` slots (16 for M <= 16, 32 otherwise) and the arrays end with 32 pad slots.
    const uint32_t *list_size;   // [nc]
    const uint32_t *list_slot;   // [nc]
    const uint8_t *codes;        // [nslots][M], byte-permuted per slot (see scan.cu)
    const uint8_t *norm_codes;   // [nslots]
    const uint8_t *sub_ids;      // [nslots] sub-group of each code (Grouping) or nullptr
    const uint32_t *ids;         // [nslots]
    // grouping metadata (nullptr for IVFADC)
    const uint32_t *nn_idx;      // [nc][nsubc]
    const uint32_t *sub_sizes;   // [nc][nsubc]
    const float *alphas;         // [nc]
    const float *icd;            // [nc][nsubc]
    const float *centroid_norms; // [nc]
    const float *centroids;      // [nc][d]  (search-time table, i.e. rotated when OPQ)
};

// Per-query work tables produced by the prep kernels, consumed by the scan kernel.  One contiguous RECORD
// of `stride` 32-bit words per query so that a slice of queries is one contiguous block (multi-GPU: ranks
// prepare disjoint query slices and all-gather the records).  Records do not depend on the shard: the scan
// kernel keeps only the lists whose centroid id lies in its own range.
//   [slot P][n P][base P][seq P][cid P][sub_base P*S]      P = nprobe, S = nsubc (0 for IVFADC)
struct SegTables {
    uint32_t *rec;        // [nq][stride]
    size_t stride;        // words per record (multiple of 4)
    int nprobe, nsubc;
    float *qsd;           // [nq][nprobe][nsubc] scratch: pruning values (K-G1 -> K-G2/K-G3), not part of the record
    float *threshold;     // [nq]
    uint32_t *nscanned;   // [nq] codes the reference rule scans (all shards)
    uint32_t *nwalked;    // [nq] codes of the ACTIVE lists (what a scan without kept-run tiling would stream; all shards)
    // PACKED records (sharded search, see launch_pack_records): query q's record starts at rec + qoff[q] and holds
    // only qnp[q] probes -- the ones the receiving rank owns -- in the same field order; null = the fixed layout
    const uint32_t *qoff = nullptr;
    const uint32_t *qnp = nullptr;
    __host__ __device__ uint32_t *slot(size_t q) const { return rec + q * stride; }                  // first slot of the list
    __host__ __device__ uint32_t *n(size_t q) const { return rec + q * stride + nprobe; }             // codes to walk (0 = skipped)
    __host__ __device__ float *base(size_t q) const { return reinterpret_cast<float *>(rec + q * stride + 2 * nprobe); }   // term1 (IVFADC)
    __host__ __device__ uint32_t *seq(size_t q) const { return rec + q * stride + 3 * nprobe; }       // scan-order number of the first code
    __host__ __device__ uint32_t *cid(size_t q) const { return rec + q * stride + 4 * nprobe; }       // centroid id (ownership test)
    __host__ __device__ float *sub_base(size_t q) const { return reinterpret_cast<float *>(rec + q * stride + 5 * nprobe); }   // term1+term2 per sub-group, +inf = pruned/empty
};
static inline size_t seg_record_words(int nprobe, int nsubc)
{
    return ((size_t)5 * nprobe + (size_t)nprobe * nsubc + 3) & ~(size_t)3;
}

// Sharded search: split the fixed-stride records of nq queries into `nparts` packed streams, one per owning rank
// (equal centroid ranges of ceil(nc / nparts)): part r receives, per query, a mini record [slot][n][base][seq][cid]
// [sub_base] of the probes whose centroid rank r owns and that the max_codes rule scans, in probe order.
//   out[pbase[r] + off[r][q] ...], cnt[r][q] probes, pwords[r] 32-bit words in part r (pbase = their exclusive scan)
cudaError_t launch_pack_records(cudaStream_t st, uint64_t *launches, int nq, int nprobe, int nsubc, int nc, int nparts,
                                const uint32_t *rec, size_t stride, uint32_t *out, uint32_t *cnt, uint32_t *off,
                                unsigned long long *pwords);
cudaError_t launch_prep_adc(cudaStream_t st, uint64_t *launches, int nq, int nprobe, uint64_t max_codes,
                            const float *coarse_d, const uint32_t *coarse_i, const DeviceLists &L, const SegTables &T);

// Workspace of the Grouping preparation (K-G1a -> K-G1b -> K-G1c), per query:
//   rows[cap]  distinct neighbour centroids: [0, nmemo) probed ones whose distance the coarse stage knows,
//              [nmemo, nmemo + nwork) rows to evaluate;   dv[cap] = (d(q, row), ||row||^2)
//   ppos[P*S]  dense row index of every (probe, sub-group) pair, 0xffff = none;   cnt = (nmemo, nwork)
struct G1Work {
    uint32_t *rows;
    float2 *dv;
    uint16_t *ppos;
    uint2 *cnt;
    int cap;          // nprobe * nsubc + nprobe
};
size_t g1_work_bytes(int nprobe, int nsubc);   // per query
cudaError_t launch_prep_grouping(cudaStream_t st, uint64_t *launches, int nq, int d, int nprobe, int nsubc,
                                 uint64_t max_codes, bool do_pruning, const float *queries,
                                 const float *coarse_d, const uint32_t *coarse_i, const DeviceLists &L,
                                 const SegTables &T, void *g1_work /* nq * g1_work_bytes() */,
                                 cudaEvent_t ev_before_g1b /* may be null */, cudaEvent_t ev_after_g1b /* may be null */,
                                 unsigned long long *rows_counter /* distinct neighbour rows, may be null */);

size_t scan_smem_bytes(int M, int nprobe, int k, int nsubc);
// partial outputs [nq][nsplit][k]; when nsplit == 1 the kernel writes the final (distances, labels) directly
// compact per-query tables lutc[q][j][m] = -2 <q_m, c_{m,j}> consumed by launch_scan
cudaError_t launch_lut(cudaStream_t st, uint64_t *launches, int M, int nq, int d, const float *queries,
                       const float *pqc, float *lutc, const uint32_t *qnp = nullptr);
cudaError_t launch_scan(cudaStream_t st, uint64_t *launches, int M, bool grouping, int nq, int d, int nprobe,
                        int nsubc, bool pruned, int k, int nsplit, const float *lutc, const float *norm_table,
                        const DeviceLists &L, const SegTables &T, uint32_t c_begin, uint32_t c_end, float *part_d,
                        uint32_t *part_seq, uint32_t *part_slot, float *out_d, int64_t *out_labels, uint32_t *out_seq);
cudaError_t launch_merge(cudaStream_t st, uint64_t *launches, int nq, int k, int nsplit, const float *part_d,
                         const uint32_t *part_seq, const uint32_t *part_slot, const uint32_t *ids,
                         float *out_d, int64_t *out_labels, uint32_t *out_seq /* optional */);
// merge of nparts shard results laid out [part][nq][k] (what an all-gather produces)
cudaError_t launch_merge_shards(cudaStream_t st, uint64_t *launches, int nq, int k, int nparts, const float *part_d,
                                const uint32_t *part_seq, const int64_t *part_labels, float *out_d,
                                int64_t *out_labels);

// ----------------------------------------------------------------------------- small dense helpers
// out[n][d] = x[n][d] * A^T (transpose == false: y = A x per row, LinearTransform::apply_noalloc)
//           = x[n][d] * A   (transpose == true : transform_transpose)
cudaError_t launch_rotate(cudaStream_t st, uint64_t *launches, size_t n, int d, const float *A, bool transpose,
                          const float *x, float *out);
// squared norms of the first dk (< 0: all d) components of every row
cudaError_t launch_row_norms(cudaStream_t st, uint64_t *launches, size_t n, int d, const float *x, float *out,
                             int dk = -1);
// dims the coarse distance really uses: hnswlib's fstdistfunc drops the d % 16 tail (hnswalg.cpp:326-356)
static inline int coarse_dims(int d) { return (d >> 4) ? ((d >> 4) << 4) : d; }
// exact re-rank of ncand candidates per query against their raw vectors (searchDisk), ascending (distance, label)
cudaError_t launch_rerank(cudaStream_t st, uint64_t *launches, int nq, int ncand, int k, int d, const float *queries,
                          const float *cand, const int64_t *cand_labels, float *out_d, int64_t *out_l);
cudaError_t launch_u8_to_f32(cudaStream_t st, uint64_t *launches, size_t n, const uint8_t *x, float *out);

// ----------------------------------------------------------------------------- coarse stage
struct CoarseWork {
    float *grp_val;     // [nq][ngroups][COARSE_T]
    uint32_t *grp_idx;  // [nq][ngroups][COARSE_T]
    uint32_t *cand;     // [nq][cand_cap] candidate centroid ids
    float *cand_d;      // [nq][cand_cap]
    uint32_t *err;      // [1] sticky overflow flag
    int cand_cap;
};
size_t coarse_work_bytes(int nq, int nc, int kprime, int *cand_cap);
// operands of the tensor-core GEMM (coarse_tc.cu): fp16 hi/lo splits of the centroids and of the
// current query batch, row pitch dp halves (a multiple of 8)
struct TcSide {
    __half *chi, *clo;      // [nc][dp]
    uint32_t *cmax;         // [2] bits of max |centroid component|, bits of max ||c||^2
    __half *qhi, *qlo;      // [batch][dp]
    float *qfac;            // [batch] -2 / (query scale * centroid scale)
    uint32_t *qstat;        // [2] impl 3: bits of max ||q||^2 over the batch, 1 if any query has a non-zero lo part
    int dp;
    // GEMM column j holds centroid id (j * idmul) mod nc: a fixed low-discrepancy permutation, so that centroids with
    // NEIGHBOURING ids (locality-ordered tables, see coarse_col_multiplier) never share a group of 128 columns -- the
    // epilogue keeps a few values per group and the select stage re-scans a group that may hide more
    float *cnp;             // [nc] ||c||^2 in column order
    uint32_t idmul;         // 0 = identity
};
// multiplier of the column -> centroid id map for a table of nc centroids (host)
uint32_t coarse_col_multiplier(size_t nc);
bool coarse_tc_supported(int d);
cudaError_t launch_tc_prepare_centroids(cudaStream_t st, uint64_t *launches, size_t nc, int d, const float *centroids,
                                        const float *cnorm_gemm, const TcSide &S);
// exact top-`nprobe` (ascending (dist, idx), fstdistfunc numerics) of every query among `nc` centroids.
// impl 0: fp32 SIMT tile GEMM; impl 1: tcgen05 split-fp16 GEMM, three products (needs the TcSide operands);
// impl 2: the same with two products (q.c_hi + q_lo.c_hi): the dropped q.c_lo term is bounded by
// 2^-11 ||q|| ||c|| and the select stage widens its threshold by twice that bound.
// impl 3: ONE product (q_hi.c_hi), both cross terms covered by the select margin (q_lo == 0 exactly for byte-valued
// vectors: then only the centroid term counts), values kept as order-preserving INTEGER keys (quantised value * 128 +
// column, built on the FMA pipe) and 2 kept per group when the groups are many.
cudaError_t launch_coarse(cudaStream_t st, uint64_t *launches, int impl, int nq, int nc, int d, int nprobe,
                          int slack, const float *queries, const float *centroids, const float *cnorm_gemm,
                          const TcSide &S, const CoarseWork &W, int self_off /* -1 = none */,
                          float *out_d, uint32_t *out_i, cudaEvent_t ev_after_gemm = nullptr);

// ----------------------------------------------------------------------------- indexing
// residual rows: out[i] = x[i] - base_row(i); base_row = centroid[idx[i]] or, for Grouping,
// centroid[g] + alpha[g] * (centroid[nn_idx[g][sub[i]]] - centroid[g])
cudaError_t launch_residual(cudaStream_t st, uint64_t *launches, size_t n, int d, const float *x,
                            const uint32_t *idx, const uint8_t *sub, const float *centroids, const uint32_t *nn_idx,
                            const float *alphas, int nsubc, float sign, const float *addend, float *out);
cudaError_t launch_pq_encode(cudaStream_t st, uint64_t *launches, size_t n, int d, int M, const float *pqc,
                             const float *residuals, uint8_t *codes);
cudaError_t launch_pq_decode(cudaStream_t st, uint64_t *launches, size_t n, int d, int M, const float *pqc,
                             const uint8_t *codes, float *out);
cudaError_t launch_norm_encode(cudaStream_t st, uint64_t *launches, size_t n, int d, const float *recon,
                               const float *norm_table, uint8_t *norm_codes, float *norms_out);
// icd[c][s] = L2sqr16(centroid[nn_idx[c][s]], centroid[c])
cudaError_t launch_gather_dists(cudaStream_t st, uint64_t *launches, int nrow, int per_row, int d,
                                const float *rows_x, const uint32_t *idx, const float *table, float *out);
cudaError_t launch_alpha_reduce(cudaStream_t st, uint64_t *launches, int nc, const uint32_t *grp_off,
                                const float *num, const float *den, float *alphas);
cudaError_t launch_hist(cudaStream_t st, uint64_t *launches, size_t n, const uint32_t *key, uint32_t *hist);

}  // namespace ivfb200
