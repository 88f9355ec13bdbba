// TEST INFRASTRUCTURE ONLY -- not part of the product path.
//
// C API over the *reference's own* classes (ivfhnsw::IndexIVF_HNSW,
// ivfhnsw::IndexIVF_HNSW_Grouping), compiled by oracle/Makefile against the
// sources where they lie under /root/reference (nothing is copied into this
// repo).  The resulting oracle/_ref/libivfhnsw_ref.so is what pins parity:
//   * tests/golden/*.npz are produced from it by oracle/make_golden.py
//   * oracle/ivfhnsw_port.c (the CPU restatement) is validated against it
//   * bench.py --impl reference / cpu_baseline.kind == "reference" time it
// Only tests/, __graft_entry__.smoke() and bench.py's cpu legs may load it.
//
// Every entry point forwards to one reference method (cited inline).
#include <cstdint>
#include <cstring>
#include <cstdio>
#include <vector>
#include <string>

#include <ivf-hnsw/IndexIVF_HNSW.h>
#include <ivf-hnsw/IndexIVF_HNSW_Grouping.h>
#include <faiss/Clustering.h>
#include <faiss/IndexFlat.h>

using namespace ivfhnsw;
typedef IndexIVF_HNSW::idx_t idx_t;

namespace {

// Subclasses only to reach the protected members the parity harness compares
// (IndexIVF_HNSW.h:80-81, IndexIVF_HNSW_Grouping.h:55-61).
struct RefADC : IndexIVF_HNSW {
    using IndexIVF_HNSW::IndexIVF_HNSW;
    std::vector<float> &cn() { return centroid_norms; }
};
struct RefGrp : IndexIVF_HNSW_Grouping {
    using IndexIVF_HNSW_Grouping::IndexIVF_HNSW_Grouping;
    std::vector<float> &cn() { return centroid_norms; }
    std::vector<std::vector<float>> &icd() { return inter_centroid_dists; }
};

struct Handle {
    bool grouping;
    RefADC *adc;
    RefGrp *grp;
    IndexIVF_HNSW *base() { return grouping ? (IndexIVF_HNSW *)grp : (IndexIVF_HNSW *)adc; }
};

}  // namespace

extern "C" {

// IndexIVF_HNSW.cpp:8-24 / IndexIVF_HNSW_Grouping.cpp:29-41
void *ref_create(size_t d, size_t nc, size_t code_size, size_t nsubc)
{
    Handle *h = new Handle();
    h->grouping = nsubc > 0;
    h->adc = nullptr;
    h->grp = nullptr;
    if (h->grouping) {
        h->grp = new RefGrp(d, nc, code_size, 8, nsubc);
        h->grp->do_pruning = false;
    } else {
        h->adc = new RefADC(d, nc, code_size, 8);
    }
    h->base()->do_opq = false;
    h->base()->nprobe = 1;
    h->base()->max_codes = 1000000000;
    return h;
}

void ref_destroy(void *hv)
{
    Handle *h = (Handle *)hv;
    if (h->adc) delete h->adc;
    if (h->grp) delete h->grp;
    delete h;
}

// IndexIVF_HNSW.cpp:40-65 (builds or loads the HNSW over the centroid file)
void ref_build_quantizer(void *hv, const char *path_centroids, const char *path_info,
                         const char *path_edges, size_t M, size_t efConstruction)
{
    ((Handle *)hv)->base()->build_quantizer(path_centroids, path_info, path_edges, M, efConstruction);
}

void ref_set_params(void *hv, size_t nprobe, size_t max_codes, size_t efSearch, int do_pruning)
{
    Handle *h = (Handle *)hv;
    h->base()->nprobe = nprobe;
    h->base()->max_codes = max_codes;
    h->base()->quantizer->efSearch = efSearch;
    if (h->grouping) h->grp->do_pruning = do_pruning != 0;
}

void ref_set_do_opq(void *hv, int do_opq) { ((Handle *)hv)->base()->do_opq = do_opq != 0; }

// tests/test_ivfhnsw_sift1b.cpp:54-65 -- codebooks are swapped in through files
void ref_load_pq(void *hv, const char *path_pq, const char *path_norm_pq, const char *path_opq)
{
    IndexIVF_HNSW *ix = ((Handle *)hv)->base();
    if (ix->pq) delete ix->pq;
    ix->pq = faiss::read_ProductQuantizer(path_pq);
    if (ix->norm_pq) delete ix->norm_pq;
    ix->norm_pq = faiss::read_ProductQuantizer(path_norm_pq);
    if (path_opq && path_opq[0]) {
        if (ix->opq_matrix) delete ix->opq_matrix;
        ix->opq_matrix = dynamic_cast<faiss::LinearTransform *>(faiss::read_VectorTransform(path_opq));
    }
}

void ref_save_pq(void *hv, const char *path_pq, const char *path_norm_pq, const char *path_opq)
{
    IndexIVF_HNSW *ix = ((Handle *)hv)->base();
    faiss::write_ProductQuantizer(ix->pq, path_pq);
    faiss::write_ProductQuantizer(ix->norm_pq, path_norm_pq);
    if (path_opq && path_opq[0] && ix->opq_matrix)
        faiss::write_VectorTransform(ix->opq_matrix, path_opq);
}

// IndexIVF_HNSW.cpp:536-593 / IndexIVF_HNSW_Grouping.cpp:486-618
void ref_train_pq(void *hv, size_t n, const float *x)
{
    Handle *h = (Handle *)hv;
    if (h->grouping) h->grp->train_pq(n, x); else h->adc->train_pq(n, x);
}

// IndexIVF_HNSW.cpp:68-72
void ref_assign(void *hv, size_t n, const float *x, uint32_t *labels)
{
    ((Handle *)hv)->base()->assign(n, x, labels, 1);
}

// hnswlib/hnswalg.cpp:227-234 unloaded exactly as IndexIVF_HNSW.cpp:248-259
// does: slot 0 = nearest.  Returns the number of results the graph produced.
size_t ref_coarse(void *hv, const float *q, size_t nprobe, float *dists, uint32_t *idxs)
{
    IndexIVF_HNSW *ix = ((Handle *)hv)->base();
    auto coarse = ix->quantizer->searchKnn(q, nprobe);
    size_t got = coarse.size();
    for (long i = (long)got - 1; i >= 0; i--) {
        dists[i] = coarse.top().first;
        idxs[i] = coarse.top().second;
        coarse.pop();
    }
    return got;
}

// IndexIVF_HNSW.cpp:75-138
void ref_add_batch(void *hv, size_t n, const float *x, const uint32_t *ids, const uint32_t *idx)
{
    ((Handle *)hv)->base()->add_batch(n, x, ids, idx);
}

// IndexIVF_HNSW_Grouping.cpp:43-157
void ref_add_group(void *hv, size_t g, size_t n, const float *x, const uint32_t *ids)
{
    ((Handle *)hv)->grp->add_group(g, n, x, ids);
}

// IndexIVF_HNSW.cpp:781-787, IndexIVF_HNSW_Grouping.cpp:620-631, IndexIVF_HNSW.cpp:789-800
void ref_compute_centroid_norms(void *hv) { ((Handle *)hv)->base()->compute_centroid_norms(); }
void ref_compute_inter_centroid_dists(void *hv) { ((Handle *)hv)->grp->compute_inter_centroid_dists(); }
void ref_rotate_quantizer(void *hv) { ((Handle *)hv)->base()->rotate_quantizer(); }

void ref_write(void *hv, const char *path)
{
    Handle *h = (Handle *)hv;
    if (h->grouping) h->grp->write(path); else h->adc->write(path);
}
void ref_read(void *hv, const char *path)
{
    Handle *h = (Handle *)hv;
    if (h->grouping) h->grp->read(path); else h->adc->read(path);
}

// IndexIVF_HNSW.cpp:234-296 / IndexIVF_HNSW_Grouping.cpp:188-363, looped over
// nq queries exactly like tests/test_ivfhnsw_sift1b.cpp:192-208 (sequential).
// Also returns the probed centroids (TRACE_CENTROIDS vectors are stored
// farthest-first, IndexIVF_HNSW.cpp:249-256; we flip them to nearest-first).
void ref_search(void *hv, size_t nq, size_t k, const float *x, float *distances, long *labels,
                float *coarse_dists, uint32_t *coarse_idxs)
{
    Handle *h = (Handle *)hv;
    IndexIVF_HNSW *ix = h->base();
    const size_t d = ix->d, np = ix->nprobe;
    for (size_t i = 0; i < nq; i++) {
        if (h->grouping) h->grp->search(k, x + i * d, distances + i * k, labels + i * k);
        else h->adc->search(k, x + i * d, distances + i * k, labels + i * k);
        if (coarse_dists && coarse_idxs) {
            const size_t got = ix->trace_centroid_idxs.size();
            for (size_t j = 0; j < np; j++) {
                size_t src = got - 1 - j;
                coarse_dists[i * np + j] = j < got ? ix->trace_query_centroid_dists[src] : 0.f;
                coarse_idxs[i * np + j] = j < got ? ix->trace_centroid_idxs[src] : 0;
            }
        }
    }
}

// IndexIVF_HNSW.cpp:453-492 -- coarse candidates supplied by the caller
void ref_search2(void *hv, size_t nq, size_t k, const float *x, float *distances, long *labels,
                 float *coarse_dists, uint32_t *coarse_idxs)
{
    IndexIVF_HNSW *ix = ((Handle *)hv)->base();
    const size_t d = ix->d, np = ix->nprobe;
    for (size_t i = 0; i < nq; i++)
        ix->search2(k, x + i * d, distances + i * k, labels + i * k,
                    coarse_dists + i * np, coarse_idxs + i * np);
}

// ---- accessors (plain copies of public / protected members) ---------------
size_t ref_list_size(void *hv, size_t c) { return ((Handle *)hv)->base()->ids[c].size(); }
void ref_get_list(void *hv, size_t c, uint32_t *ids, uint8_t *codes, uint8_t *norm_codes)
{
    IndexIVF_HNSW *ix = ((Handle *)hv)->base();
    memcpy(ids, ix->ids[c].data(), ix->ids[c].size() * sizeof(uint32_t));
    memcpy(codes, ix->codes[c].data(), ix->codes[c].size());
    memcpy(norm_codes, ix->norm_codes[c].data(), ix->norm_codes[c].size());
}
void ref_get_centroid(void *hv, size_t c, float *out)
{
    IndexIVF_HNSW *ix = ((Handle *)hv)->base();
    memcpy(out, ix->quantizer->getDataByInternalId(c), ix->d * sizeof(float));
}
void ref_get_centroid_norms(void *hv, float *out)
{
    Handle *h = (Handle *)hv;
    std::vector<float> &v = h->grouping ? h->grp->cn() : h->adc->cn();
    memcpy(out, v.data(), v.size() * sizeof(float));
}
void ref_get_pq(void *hv, float *pq_centroids, float *norm_centroids)
{
    IndexIVF_HNSW *ix = ((Handle *)hv)->base();
    memcpy(pq_centroids, ix->pq->centroids.data(), ix->pq->centroids.size() * sizeof(float));
    memcpy(norm_centroids, ix->norm_pq->centroids.data(), ix->norm_pq->centroids.size() * sizeof(float));
}
int ref_get_opq(void *hv, float *A)
{
    IndexIVF_HNSW *ix = ((Handle *)hv)->base();
    if (!ix->opq_matrix) return 0;
    memcpy(A, ix->opq_matrix->A.data(), ix->opq_matrix->A.size() * sizeof(float));
    return 1;
}
// grouping metadata; returns 0 when the group was never added (vectors empty)
int ref_get_group_meta(void *hv, size_t c, uint32_t *nn_idxs, uint32_t *sub_sizes, float *alpha, float *icd)
{
    RefGrp *g = ((Handle *)hv)->grp;
    const size_t nsubc = g->nsubc;
    int full = 1;
    if (g->nn_centroid_idxs[c].size() == nsubc) memcpy(nn_idxs, g->nn_centroid_idxs[c].data(), nsubc * 4);
    else full = 0;
    if (g->subgroup_sizes[c].size() == nsubc) memcpy(sub_sizes, g->subgroup_sizes[c].data(), nsubc * 4);
    else { memset(sub_sizes, 0, nsubc * 4); }
    *alpha = g->alphas[c];
    if (g->icd()[c].size() == nsubc) memcpy(icd, g->icd()[c].data(), nsubc * 4);
    return full;
}

// ivfhnsw::fvec_L2sqr (utils.cpp:22-51), for unit checks of the port
float ref_fvec_L2sqr(const float *x, const float *y, size_t d) { return ivfhnsw::fvec_L2sqr(x, y, d); }

}  // extern "C"

// faiss::Clustering::train over an IndexFlatL2 (faiss/Clustering.cpp:67-230): the k-means the reference's tooling
// would use for the centroid file; returns the last iteration's objective.
extern "C" double ref_kmeans(size_t d, size_t n, size_t k, const float *x, int niter, int seed, float *centroids_out)
{
    faiss::Clustering clus((int)d, (int)k);
    clus.niter = niter;
    clus.seed = seed;
    clus.verbose = false;
    clus.max_points_per_centroid = 1 << 30;     // no sub-sampling: same training set on both sides
    faiss::IndexFlatL2 index(d);
    clus.train(n, x, index);
    memcpy(centroids_out, clus.centroids.data(), sizeof(float) * d * k);
    return clus.obj.empty() ? 0.0 : (double)clus.obj.back();
}
