// <ivf-hnsw/IndexIVF_HNSW_Grouping.h> of the B200 drop-in (reference IndexIVF_HNSW_Grouping.h:14-75).
#pragma once
#include "IndexIVF_HNSW.h"

namespace ivfhnsw {

struct IndexIVF_HNSW_Grouping : IndexIVF_HNSW {
    size_t nsubc;      // sub-centroids per group
    bool do_pruning;

    std::vector<std::vector<idx_t> > nn_centroid_idxs;   // host mirrors, filled by sync_host_lists()
    std::vector<std::vector<idx_t> > subgroup_sizes;
    std::vector<float> alphas;

public:
    IndexIVF_HNSW_Grouping(size_t dim, size_t ncentroids, size_t bytes_per_code, size_t nbits_per_idx,
                           size_t nsubcentroids);

    // thread-safe for distinct group_idx, as the reference's OpenMP ingest loop requires
    // (tests/test_ivfhnsw_grouping_sift1b.cpp:177-194): points are staged under a mutex and flushed to the GPU
    void add_group(size_t group_idx, size_t group_size, const float *x, const idx_t *ids);
    void search(size_t k, const float *x, float *distances, long *labels);
    // ANN search for 2k candidates, then exact re-rank against the raw vectors of the .bvecs base file
    // (reference IndexIVF_HNSW_Grouping.h:40, .cpp:365-395).  The candidate search and the re-rank run on the GPU
    // (ivfhnsw_b200_search + ivfhnsw_b200_rerank); only the 2k vector reads touch the file.
    void searchDisk(size_t k, const float *query, float *distances, long *labels, const char *path_base);
    // nq queries at once (B200 extension): rows sorted ascending by (exact distance, label)
    void searchDisk_batch(size_t nq, size_t k, const float *queries, float *distances, long *labels, const char *path_base);
    void dump_inter_centroid_dists(char *path);     // text dump, reference .cpp:633-653
    void write(const char *path_index);
    void write(const char *path_index, bool do_trunc);
    void read(const char *path_index);
    void train_pq(size_t n, const float *x);
    void compute_inter_centroid_dists();
    void sync_host_lists();
};

}  // namespace ivfhnsw
