/* ivfhnsw_b200.h -- C ABI of the B200-native IVF-HNSW search + indexing path.
 *
 * This is the drop-in boundary.  The reference has no FFI
 * layer: its boundary is the C++ API of ivfhnsw::IndexIVF_HNSW /
 * IndexIVF_HNSW_Grouping.  The C++ shims in include/ivf-hnsw/ forward every hot
 * method to one entry point below; the reference method each one replaces is
 * cited next to it (paths relative to the reference root).
 *
 * Conventions: every function returns 0 on success and a negative value on
 * failure (message via ivfhnsw_b200_last_error); nothing throws.  Pointers are
 * HOST pointers unless the parameter name ends in _dev.  The library needs an
 * sm_100a device: there is NO CPU fallback -- without a usable GPU every
 * compute call fails with IVFHNSW_B200_ENODEV.
 *
 * Result layout of the search calls: distances/labels are sorted ascending by
 * (distance, scan order); unfilled slots hold FLT_MAX / -1 exactly like
 * faiss::maxheap_heapify (faiss/Heap.h:224-233).  The C++ shim re-orders them
 * into the reference's max-heap layout.
 */
#ifndef IVFHNSW_B200_H
#define IVFHNSW_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ivfhnsw_b200 ivfhnsw_b200_t;

#define IVFHNSW_B200_OK 0
#define IVFHNSW_B200_EINVAL (-1)  /* bad argument / unsupported shape           */
#define IVFHNSW_B200_ECUDA (-2)   /* a CUDA call or kernel failed               */
#define IVFHNSW_B200_ENODEV (-3)  /* no usable sm_100 device                    */
#define IVFHNSW_B200_ESTATE (-4)  /* call order violated (e.g. search w/o lists)*/

/* ABI version of this header; ivfhnsw_b200_abi_version() must return it. */
#define IVFHNSW_B200_ABI 1
int ivfhnsw_b200_abi_version(void);

/* -- lifetime ---------------------------------------------------------------
 * IndexIVF_HNSW::IndexIVF_HNSW (IndexIVF_HNSW.cpp:8-24) and
 * IndexIVF_HNSW_Grouping::IndexIVF_HNSW_Grouping (IndexIVF_HNSW_Grouping.cpp:29-41).
 * nsubc == 0 selects IVFADC, nsubc > 0 Grouping (nsubc <= 256).
 * code_size (= PQ M) must be 8, 16, 32 or 64 and divide d; nbits must be 8. */
int ivfhnsw_b200_create(ivfhnsw_b200_t **out, size_t d, size_t nc, size_t code_size, size_t nbits,
                        size_t nsubc, int device);
void ivfhnsw_b200_destroy(ivfhnsw_b200_t *h);
const char *ivfhnsw_b200_last_error(const ivfhnsw_b200_t *h);

/* -- sharding: this handle owns the inverted lists of
 * centroids [c_begin, c_end) only; all metadata stays replicated.  Lists
 * outside the range still count for the max_codes rule but are not scanned.
 * Default is [0, nc).
 * Called BEFORE the lists are built (finalize / load_lists / read_index), the
 * build keeps only the owned lists' codes and ids in HBM (1/N of the index per
 * rank); records of other lists are counted in the size tables and dropped.
 * Called afterwards it only narrows what is scanned and must stay inside the
 * resident range (ESTATE otherwise).  A handle whose resident range is not
 * [0, nc) cannot export_lists / write_index / take further add_batch calls. */
int ivfhnsw_b200_set_shard(ivfhnsw_b200_t *h, size_t c_begin, size_t c_end);
/* Sharded CONSTRUCTION: each rank set_shard()s, ingests only the points of its own lists (add_batch with given
 * assignments, add_group for its groups), finalizes, then the ranks SUM these per-list tables (their supports are
 * disjoint: an all-reduce) and install the sums, after which every rank prunes / applies max_codes exactly like a
 * rank holding the whole index.  Device pointers: list_size [nc], sub_sizes [nc*nsubc], alphas [nc] (the last two
 * NULL / ignored without Grouping).  set checks that the resident lists keep their sizes.  No counterpart in the
 * reference (single-process). */
int ivfhnsw_b200_get_list_meta_dev(ivfhnsw_b200_t *h, uint32_t *list_size_dev, uint32_t *sub_sizes_dev, float *alphas_dev);
int ivfhnsw_b200_set_list_meta_dev(ivfhnsw_b200_t *h, const uint32_t *list_size_dev, const uint32_t *sub_sizes_dev,
                                   const float *alphas_dev);

/* -- model state ------------------------------------------------------------*/
/* centroid table = hnswlib::HierarchicalNSW::getDataByInternalId (hnswlib/hnswalg.h:76-78),
 * filled by IndexIVF_HNSW::build_quantizer (IndexIVF_HNSW.cpp:40-65); nc x d row-major */
int ivfhnsw_b200_set_centroids(ivfhnsw_b200_t *h, const float *centroids);
/* faiss::ProductQuantizer::centroids of index->pq, layout [M][256][dsub] (faiss/ProductQuantizer.h) */
int ivfhnsw_b200_set_pq(ivfhnsw_b200_t *h, const float *pq_centroids);
/* index->norm_pq centroids: 256 floats (IndexIVF_HNSW.cpp:14) */
int ivfhnsw_b200_set_norm_pq(ivfhnsw_b200_t *h, const float *norm_centroids);
/* index->opq_matrix->A, d x d row-major [d_out][d_in] (faiss/VectorTransform.h:84-85); NULL clears */
int ivfhnsw_b200_set_opq(ivfhnsw_b200_t *h, const float *A);
/* public members nprobe, max_codes, do_opq, do_pruning (IndexIVF_HNSW.h:58-62, IndexIVF_HNSW_Grouping.h:18) */
int ivfhnsw_b200_set_params(ivfhnsw_b200_t *h, size_t nprobe, size_t max_codes, int do_opq, int do_pruning);

/* IndexIVF_HNSW::compute_centroid_norms (IndexIVF_HNSW.cpp:781-787) */
int ivfhnsw_b200_compute_centroid_norms(ivfhnsw_b200_t *h);
int ivfhnsw_b200_get_centroid_norms(ivfhnsw_b200_t *h, float *out /* nc */);
int ivfhnsw_b200_set_centroid_norms(ivfhnsw_b200_t *h, const float *in /* nc, e.g. from read() */);
/* IndexIVF_HNSW::rotate_quantizer (IndexIVF_HNSW.cpp:789-800): centroids <- centroids * A^T on the device */
int ivfhnsw_b200_rotate_quantizer(ivfhnsw_b200_t *h);
int ivfhnsw_b200_get_centroids(ivfhnsw_b200_t *h, float *out /* nc*d */);

/* -- training ---------------------------------------------------------------*/
/* IndexIVF_HNSW::train_pq / IndexIVF_HNSW_Grouping::train_pq (IndexIVF_HNSW.cpp:536-593,
 * IndexIVF_HNSW_Grouping.cpp:486-618): assign, residuals (to the sub-centroid for Grouping), OPQ rotation
 * when do_opq (OPQMatrix::train, faiss/VectorTransform.cpp:587-744), M x 256-way k-means
 * (faiss/Clustering.cpp:67-230, 25 iterations), then the 1-D norm codebook.  n x d host vectors. */
int ivfhnsw_b200_train_pq(ivfhnsw_b200_t *h, size_t n, const float *x);
/* host copies of the model for index->pq->centroids / norm_pq->centroids / opq_matrix->A (any may be NULL) */
int ivfhnsw_b200_get_pq(ivfhnsw_b200_t *h, float *pq_centroids, float *norm_centroids, float *opq_A);

/* Coarse k-means for the centroid file.  The reference does not train its centroids ("centroids file
 * create by k-means tool, not here", IndexIVF_HNSW.cpp:703); this is faiss::Clustering::train (faiss/Clustering.cpp:
 * 67-230) with K-A1 as the assignment step and km_update_centroids (faiss/utils.cpp:1401-1484, void clusters re-seeded
 * with faiss' random stream) as the update.  x_dev [n][d] float32 DEVICE vectors, n >= nc; the result becomes the
 * handle's centroid table (read it back with get_centroids).  objective_out = sum of squared distances of the last
 * assignment, nsplit_out = void clusters re-seeded over all iterations (either may be NULL). */
int ivfhnsw_b200_train_centroids(ivfhnsw_b200_t *h, size_t n, const float *x_dev, int niter, long seed,
                                 double *objective_out, size_t *nsplit_out);

/* -- indexing ---------------------------------------------------------------*/
/* IndexIVF_HNSW::assign (IndexIVF_HNSW.cpp:68-72): labels[i] = searchKnn(x_i, k).top() = the k-th nearest centroid,
 * exact (k = 1 in every caller of the reference: the nearest; 1 <= k <= min(nc, 2048)) */
int ivfhnsw_b200_assign(ivfhnsw_b200_t *h, size_t n, const float *x, uint32_t *labels, size_t k);
/* same for device-resident vectors (asynchronous on the handle's stream; base vectors are NOT rotated: the OPQ
 * rotation applies to residuals, IndexIVF_HNSW.cpp:85-103) */
int ivfhnsw_b200_assign_dev(ivfhnsw_b200_t *h, size_t n, const float *x_dev, uint32_t *labels_dev);
/* same, base vectors given as bytes (what readXvecFvec<uint8_t> expands, utils.h:112-127) */
int ivfhnsw_b200_assign_u8(ivfhnsw_b200_t *h, size_t n, const uint8_t *x, uint32_t *labels);
/* IndexIVF_HNSW::add_batch (IndexIVF_HNSW.cpp:75-138); precomputed_idx may be NULL (assign first) */
int ivfhnsw_b200_add_batch(ivfhnsw_b200_t *h, size_t n, const float *x, const uint32_t *xids,
                           const uint32_t *precomputed_idx);
/* IndexIVF_HNSW::add_batch2 (IndexIVF_HNSW.cpp:140-204): add_batch that also hands the residual codes and norm codes
 * of the batch back (the caller packs them into its ORCV output buffer); codes_out [n][code_size], norm_codes_out [n] */
int ivfhnsw_b200_add_batch_codes(ivfhnsw_b200_t *h, size_t n, const float *x, const uint32_t *xids,
                                 const uint32_t *idx, uint8_t *codes_out, uint8_t *norm_codes_out);
/* IndexIVF_HNSW_Grouping::add_group (IndexIVF_HNSW_Grouping.cpp:43-157); thread-safe per distinct group.
 * STREAMED like the reference's bounded pass (tests/test_ivfhnsw_grouping_sift1b.cpp:138-201): one call delivers a
 * complete group, the raw vectors are staged in HBM and -- as soon as "stage_bytes" of them wait (set_option,
 * default 4 GiB) -- alpha, sub-groups and codes of the staged groups are computed and the raw vectors dropped;
 * finalize() handles the remainder.  A group may therefore be added ONCE (a second add_group of an already
 * encoded group fails with ESTATE), which is also the reference's contract. */
int ivfhnsw_b200_add_group(ivfhnsw_b200_t *h, size_t group_idx, size_t group_size, const float *x,
                           const uint32_t *ids);
/* device-resident bulk variant used by the ingest path: x_dev [n][d] f32, ids_dev [n], idx_dev [n] list of every
 * point or NULL (= assigned here, K-A1); any order, the per-list order of appearance is kept.  IVFADC: encoded at
 * once.  Grouping: the raw vectors are staged until ivfhnsw_b200_flush_groups()/finalize() -- the caller delivers
 * the points of a RANGE of groups (any number of calls), then flushes: that is the reference's groups_per_iter pass. */
int ivfhnsw_b200_add_points_dev(ivfhnsw_b200_t *h, size_t n, const float *x_dev, const uint32_t *ids_dev,
                                const uint32_t *idx_dev);
/* Grouping only: every group that has staged raw points is COMPLETE -- compute its alpha (compute_alpha, :691-728),
 * sub-group ids (:673-689) and residual codes (:101-131) now and release the raw vectors.  Peak HBM of an ingest is
 * bounded by the largest flush, not by the base set (1B x 128: 21 GB of records instead of 512 GB of vectors). */
int ivfhnsw_b200_flush_groups(ivfhnsw_b200_t *h);
/* flush point: builds the device CSR lists (stable by list [, sub-group]) from everything added so far.
 * For Grouping also computes nn_centroid_idxs / alphas / sub-groups (add_group body) and
 * compute_inter_centroid_dists (IndexIVF_HNSW_Grouping.cpp:620-631). */
int ivfhnsw_b200_finalize(ivfhnsw_b200_t *h);

/* -- list import / export: the per-list vectors of read()/write()
 * (IndexIVF_HNSW.cpp:637-663,758-779; IndexIVF_HNSW_Grouping.cpp:397-483), flattened as CSR.
 * list_off[nc+1]; ids[N]; codes[N*code_size]; norm_codes[N];
 * Grouping only (else NULL): nn_idx[nc*nsubc], sub_sizes[nc*nsubc], alphas[nc], icd[nc*nsubc]. */
int ivfhnsw_b200_load_lists(ivfhnsw_b200_t *h, const int64_t *list_off, const uint32_t *ids,
                            const uint8_t *codes, const uint8_t *norm_codes, const uint32_t *nn_idx,
                            const uint32_t *sub_sizes, const float *alphas, const float *icd);
int ivfhnsw_b200_ntotal(ivfhnsw_b200_t *h, size_t *ntotal);
/* ids[c].size() of every list (what the reference's tracing prints, IndexIVF_HNSW.cpp:298-326); sizes [nc] */
int ivfhnsw_b200_list_sizes(ivfhnsw_b200_t *h, uint32_t *sizes);
int ivfhnsw_b200_export_lists(ivfhnsw_b200_t *h, int64_t *list_off, uint32_t *ids, uint8_t *codes,
                              uint8_t *norm_codes, uint32_t *nn_idx, uint32_t *sub_sizes, float *alphas,
                              float *icd);

/* IndexIVF_HNSW::write / read and IndexIVF_HNSW_Grouping::write / read (IndexIVF_HNSW.cpp:637-663,758-779;
 * IndexIVF_HNSW_Grouping.cpp:397-483): the reference's own .index file, byte for byte -- the reference reads what
 * this writes and vice versa.  write finalizes first; read needs a handle created with the file's (d, nc, nsubc). */
int ivfhnsw_b200_write_index(ivfhnsw_b200_t *h, const char *path);
int ivfhnsw_b200_read_index(ivfhnsw_b200_t *h, const char *path);

/* -- search -----------------------------------------------------------------*/
/* coarse stage alone = quantizer->searchKnn(query, nprobe) unloaded nearest-first
 * (IndexIVF_HNSW.cpp:248-259), but EXACT.  x are raw queries (OPQ applied inside when do_opq). */
int ivfhnsw_b200_coarse(ivfhnsw_b200_t *h, size_t nq, const float *x, size_t nprobe, float *dists,
                        uint32_t *idxs);
/* IndexIVF_HNSW::search / IndexIVF_HNSW_Grouping::search (IndexIVF_HNSW.cpp:234-296,
 * IndexIVF_HNSW_Grouping.cpp:188-363) for nq queries at once */
int ivfhnsw_b200_search(ivfhnsw_b200_t *h, size_t nq, size_t k, const float *x, float *distances,
                        int64_t *labels);
/* IndexIVF_HNSW::search2 (IndexIVF_HNSW.cpp:453-492): coarse candidates supplied by the caller
 * ([nq][nprobe], nearest first).  Also accepted for Grouping handles (parity seam). */
int ivfhnsw_b200_search_given_coarse(ivfhnsw_b200_t *h, size_t nq, size_t k, const float *x,
                                     const float *coarse_dists, const uint32_t *coarse_idxs,
                                     float *distances, int64_t *labels);
/* device-resident variants (inputs/outputs already in HBM, asynchronous on the handle's stream
 * until ivfhnsw_b200_sync) -- what bench.py times for the "value" figure */
int ivfhnsw_b200_search_dev(ivfhnsw_b200_t *h, size_t nq, size_t k, const float *x_dev, float *distances_dev,
                            int64_t *labels_dev);
int ivfhnsw_b200_search_given_coarse_dev(ivfhnsw_b200_t *h, size_t nq, size_t k, const float *x_dev,
                                         const float *coarse_dists_dev, const uint32_t *coarse_idxs_dev,
                                         float *distances_dev, int64_t *labels_dev);
int ivfhnsw_b200_coarse_dev(ivfhnsw_b200_t *h, size_t nq, const float *x_dev, size_t nprobe,
                            float *dists_dev, uint32_t *idxs_dev);
/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).  The exact coarse
 * search of ivfhnsw_b200_coarse_dev (fp32 GEMM + select + fstdistfunc re-rank; same (dist, idx) order and values),
 * with the GEMM's intermediates in ONE file-scope device array of IVFHNSW_B200_NOALLOC_SCRATCH_BYTES that every
 * GPU loading the library holds until exit: the call allocates nothing (no cudaMalloc/cudaFree, no allocating
 * library routine) and only enqueues on the handle's stream.  xq_dev [nq][d] are queries in the space of the
 * centroid table (already rotated when OPQ is used: no rotation is applied here).  Limit: nq * ceil(nc/128) * 32
 * + nc * 4 + 768 <= IVFHNSW_B200_NOALLOC_SCRATCH_BYTES, else EINVAL (e.g. 512 queries at nc = 65536).
 * flags_dev (may be NULL) receives the select stage's overflow flags, 0 = the result is exact.  All calls on a
 * device share the array: calls that overlap on the GPU (other streams, other handles) race; serialise them. */
#define IVFHNSW_B200_NOALLOC_SCRATCH_BYTES (8u << 20)
int ivfhnsw_b200_coarse_noalloc_dev(ivfhnsw_b200_t *h, size_t nq, const float *xq_dev, size_t nprobe,
                                    float *dists_dev, uint32_t *idxs_dev, uint32_t *flags_dev);
/* sharded search: this handle scans only the lists it owns (set_shard) and returns its
 * sorted partial top-k together with the global scan-order key of every hit; coarse candidates NULL =
 * computed here over the (replicated) centroid table, identical on every rank */
int ivfhnsw_b200_search_partial_dev(ivfhnsw_b200_t *h, size_t nq, size_t k, const float *x_dev,
                                    const float *coarse_dists_dev, const uint32_t *coarse_idxs_dev,
                                    float *part_d_dev, uint32_t *part_seq_dev, int64_t *part_labels_dev);
/* merge the partial results of nparts shards laid out [part][nq][k] (the output of an NCCL all-gather)
 * into the final top-k by (distance, scan order) -- the same order a single GPU produces */
int ivfhnsw_b200_merge_partials_dev(ivfhnsw_b200_t *h, size_t nq, size_t k, size_t nparts,
                                    const float *part_d_dev, const uint32_t *part_seq_dev,
                                    const int64_t *part_labels_dev, float *distances_dev, int64_t *labels_dev);
/* Two-stage form of search_partial_dev for query-parallel preparation: the work tables
 * (coarse probes, Grouping thresholds and sub-group bases, max_codes cut) do not depend on the shard, so each
 * rank prepares a slice [q_begin, q_end) of the nq queries into records q_begin.. of tables_dev
 * (nq * tables_bytes, one contiguous record per query), the ranks all-gather the records, and every rank
 * scans its own lists for all queries. */
int ivfhnsw_b200_tables_bytes(ivfhnsw_b200_t *h, size_t *bytes_per_query);
int ivfhnsw_b200_prepare_dev(ivfhnsw_b200_t *h, size_t nq, const float *x_dev, size_t q_begin, size_t q_end,
                             void *tables_dev);
int ivfhnsw_b200_scan_tables_dev(ivfhnsw_b200_t *h, size_t nq, size_t k, const float *x_dev, const void *tables_dev,
                                 float *part_d_dev, uint32_t *part_seq_dev, int64_t *part_labels_dev);
/* The re-rank half of IndexIVF_HNSW_Grouping::searchDisk (IndexIVF_HNSW_Grouping.cpp:365-395; getL2Distance +
 * fvec_L2sqr, utils.cpp:22-51,98-137): exact distances of ncand candidate RAW vectors per query (cand [nq][ncand][d]
 * floats as readXvecFvec<uint8_t> delivers them, cand_labels [nq][ncand], < 0 = padding), the k best per query
 * ascending by (distance, label).  Host pointers. */
int ivfhnsw_b200_rerank(ivfhnsw_b200_t *h, size_t nq, size_t ncand, size_t k, const float *queries, const float *cand,
                        const int64_t *cand_labels, float *distances, int64_t *labels);
/* Sharded search without the all-gather of whole records: the HOME rank of a query (the one that prepared it) splits
 * its records into one PACKED stream per owning rank -- per query a mini record of only the probes that rank owns and
 * the max_codes rule scans -- which the ranks exchange with ONE all-to-all-v (~(5 + nsubc) words per probe instead of
 * the whole 35 KB record to everybody).  tables_dev: nq records as written by prepare_dev.  Outputs: packed_dev (capacity
 * nq * tables_bytes) = the nparts streams back to back, cnt_dev / off_dev [nparts][nq] = probes and 32-bit-word offset
 * (inside its stream) of every (part, query), part_words_dev [nparts] = length of every stream in 32-bit words.
 * Ownership = equal centroid ranges of ceil(nc / nparts), the same split as set_shard in an nparts-rank deployment. */
int ivfhnsw_b200_pack_records_dev(ivfhnsw_b200_t *h, size_t nq, const void *tables_dev, size_t nparts, void *packed_dev,
                                  uint32_t *cnt_dev, uint32_t *off_dev, uint64_t *part_words_dev);
/* scan_tables_dev over packed records: query q's record starts at 32-bit word qoff_dev[q] of packed_dev and holds
 * qnp_dev[q] probes (0 = nothing of this query lives here: an empty partial result is written) */
int ivfhnsw_b200_scan_packed_dev(ivfhnsw_b200_t *h, size_t nq, size_t k, const float *x_dev, const void *packed_dev,
                                 const uint32_t *qoff_dev, const uint32_t *qnp_dev, float *part_d_dev,
                                 uint32_t *part_seq_dev, int64_t *part_labels_dev);
int ivfhnsw_b200_sync(ivfhnsw_b200_t *h);
/* run on a caller-owned CUDA stream (cudaStream_t, e.g. torch's current stream); NULL = back to a private one.
 * STREAM ORDERING: every *_dev entry point only ENQUEUES work on the handle's stream.  Device buffers
 * produced or consumed on another stream must be ordered by the caller (an event, or ivfhnsw_b200_sync()
 * after the call and a synchronize of the producing stream before it).  The host-pointer entry points
 * (search, assign, add_batch, ...) synchronise internally and need nothing. */
int ivfhnsw_b200_set_stream(ivfhnsw_b200_t *h, void *stream);

/* -- introspection for tests, bench and profiles ----------------------------*/
typedef struct ivfhnsw_b200_stats {
    uint64_t kernel_launches;   /* kernels of this library launched since create/reset          */
    uint64_t scanned_codes;     /* codes whose ADC distance was required by the reference rule  */
    uint64_t walked_codes;      /* codes of the active lists, pruned sub-groups included (>= scanned) */
    uint64_t queries;           /* queries searched                                             */
    double scan_ms;             /* CUDA-event time spent in the ADC scan kernel                 */
    double coarse_ms;           /* ... in the coarse GEMM + select kernels                      */
    double prep_ms;             /* ... in work-list / sub-centroid kernels                      */
    uint64_t scan_launches;     /* number of scan-kernel launches included in scan_ms           */
    uint64_t coarse_flops;      /* 2*nc*d per query/vector pushed through the coarse GEMM       */
    double gemm_ms;             /* ... of coarse_ms, the tensor-core GEMM kernel alone          */
    double g1_ms;               /* ... of prep_ms, the sub-centroid distance kernel (K-G1b)     */
    double lut_ms;              /* ... of scan_ms, the lookup-table kernel (K-L1)               */
    uint64_t g1_rows;           /* distinct neighbour-centroid rows K-G1 gathered               */
    uint64_t prep_launches;     /* number of prep batches included in coarse_ms / prep_ms       */
} ivfhnsw_b200_stats_t;
/* timing != 0 brackets each stage with CUDA events on the handle's stream (adds a sync per call) */
int ivfhnsw_b200_set_timing(ivfhnsw_b200_t *h, int timing);
int ivfhnsw_b200_get_stats(ivfhnsw_b200_t *h, ivfhnsw_b200_stats_t *out);
int ivfhnsw_b200_reset_stats(ivfhnsw_b200_t *h);
/* Grouping debug outputs of the last search call's first min(nq, cap) queries: pruning threshold per
 * query (IndexIVF_HNSW_Grouping.cpp:261) and codes scanned per query */
int ivfhnsw_b200_last_thresholds(ivfhnsw_b200_t *h, size_t cap, float *thresholds, uint32_t *scanned);
/* tuning knobs: "coarse_slack", "coarse_impl" (0 = SIMT fp32, 1 = tcgen05 split-fp16 three products, 2 = two products
 * with a widened select threshold, 3 = one product + integer keys, the default), "scan_split" (0 = auto), "stage_bytes" (raw Grouping vectors staged
 * before add_group encodes them) */
int ivfhnsw_b200_set_option(ivfhnsw_b200_t *h, const char *name, long value);

#ifdef __cplusplus
}
#endif
#endif /* IVFHNSW_B200_H */
