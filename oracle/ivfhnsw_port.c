/* TEST INFRASTRUCTURE ONLY -- CPU restatement ("port") of the uniio/ivf-hnsw
 * search + indexing hot path, in plain C, one thread, no SIMD, no BLAS.
 *
 * It exists so that tests/ can check the CUDA path on a box that has no
 * /root/reference.  It is NOT the product and NOT a fallback: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load it; ivf-hnsw_b200 never does.
 *
 * Parity status: PINNED.  The reference ships no golden vectors,
 * so the pin is made from the compiled reference itself: oracle/_ref/
 * libivfhnsw_ref.so (oracle/Makefile) -> oracle/make_golden.py ->
 * tests/golden/*.npz, and tests/test_oracle.py checks every function below
 * against those fixtures (and against the live _ref library when present).
 *
 * All arithmetic is fp32 like the reference.  Summation orders mimic the
 * reference's AVX/SSE lane structure (cited per function); since the reference
 * itself is built with -Ofast (CMakeLists.txt:22) its last-ulp behaviour is
 * compiler-chosen, so comparisons use 1e-5..1e-4 relative tolerances, never
 * bit equality, for floating-point outputs.  Integer outputs (labels, codes,
 * sub-group ids) are compared exactly away from fp ties.
 *
 * Index storage is passed as flat CSR arrays (what the .index file holds,
 * IndexIVF_HNSW.cpp:637-663 / IndexIVF_HNSW_Grouping.cpp:397-438):
 *   list_off[nc+1] (int64), ids[N] u32, codes[N*code_size] u8, norm_codes[N] u8
 *   grouping: nn_idx[nc*nsubc] u32, sub_sizes[nc*nsubc] u32, alphas[nc],
 *             icd[nc*nsubc] (inter_centroid_dists), centroid_norms[nc]
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>
#include <math.h>

#define PORT_EPS 0.00001f /* utils.h:31 */

/* ------------------------------------------------------------------ kernels */

/* ivfhnsw::fvec_L2sqr, utils.cpp:22-51 == hnswlib fstdistfunc,
 * hnswlib/hnswalg.cpp:326-356: only floor(d/16)*16 dims, 8 lane accumulators,
 * lanes summed left to right. */
float port_fvec_L2sqr16(const float *x, const float *y, size_t d)
{
    float lane[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    size_t n = (d >> 4) << 4;
    for (size_t i = 0; i < n; i += 8)
        for (int l = 0; l < 8; l++) {
            float df = x[i + l] - y[i + l];
            lane[l] = lane[l] + df * df;
        }
    float r = lane[0];
    for (int l = 1; l < 8; l++) r = r + lane[l];
    return r;
}

/* helper: finish of the faiss AVX kernels (faiss/utils.cpp:514-531,548-567):
 * fold 8 lanes to 4 (hi + lo), then optional 4-chunk, masked tail, 2x hadd */
static float fold_tail(float *l8, const float *x, const float *y, size_t d, size_t i, int mode)
{
    float l4[4];
    for (int l = 0; l < 4; l++) l4[l] = l8[l + 4] + l8[l];
    if (d - i >= 4) {
        for (int l = 0; l < 4; l++) {
            float a = x[i + l], b = y ? y[i + l] : 0.f;
            float t = mode == 0 ? a * b : (a - b) * (a - b);
            l4[l] = l4[l] + t;
        }
        i += 4;
    }
    for (int l = 0; i + l < d; l++) {
        float a = x[i + l], b = y[i + l];
        float t = mode == 0 ? a * b : (a - b) * (a - b);
        l4[l] = l4[l] + t;
    }
    return (l4[0] + l4[1]) + (l4[2] + l4[3]);
}

/* faiss::fvec_inner_product, faiss/utils.cpp:500-532 */
float port_fvec_inner_product(const float *x, const float *y, size_t d)
{
    float l8[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    size_t i = 0;
    for (; i + 8 <= d; i += 8)
        for (int l = 0; l < 8; l++) l8[l] = l8[l] + x[i + l] * y[i + l];
    return fold_tail(l8, x, y, d, i, 0);
}

/* faiss::fvec_L2sqr (all d dims), faiss/utils.cpp:534-568 */
float port_faiss_L2sqr(const float *x, const float *y, size_t d)
{
    float l8[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    size_t i = 0;
    for (; i + 8 <= d; i += 8)
        for (int l = 0; l < 8; l++) {
            float df = x[i + l] - y[i + l];
            l8[l] = l8[l] + df * df;
        }
    return fold_tail(l8, x, y, d, i, 1);
}

/* faiss::fvec_norm_L2sqr, faiss/utils.cpp:632-650: 4 SSE lanes + masked tail */
float port_fvec_norm_L2sqr(const float *x, size_t d)
{
    float l4[4] = {0, 0, 0, 0};
    size_t i = 0;
    for (; i + 4 <= d; i += 4)
        for (int l = 0; l < 4; l++) l4[l] = l4[l] + x[i + l] * x[i + l];
    for (int l = 0; i + l < d; l++) l4[l] = l4[l] + x[i + l] * x[i + l];
    return (l4[0] + l4[1]) + (l4[2] + l4[3]);
}

/* LinearTransform::apply_noalloc without bias, faiss/VectorTransform.cpp:106-132:
 * xt[i][o] = sum_j A[o][j] * x[i][j]  (A row-major [d_out][d_in]) */
void port_opq_apply(size_t n, size_t d, const float *A, const float *x, float *xt)
{
    for (size_t i = 0; i < n; i++)
        for (size_t o = 0; o < d; o++) {
            float s = 0.f;
            for (size_t j = 0; j < d; j++) s += A[o * d + j] * x[i * d + j];
            xt[i * d + o] = s;
        }
}

/* LinearTransform::transform_transpose, faiss/VectorTransform.cpp:135-158:
 * x[i][j] = sum_o A[o][j] * y[i][o] */
void port_opq_apply_transpose(size_t n, size_t d, const float *A, const float *y, float *x)
{
    for (size_t i = 0; i < n; i++)
        for (size_t j = 0; j < d; j++) {
            float s = 0.f;
            for (size_t o = 0; o < d; o++) s += A[o * d + j] * y[i * d + o];
            x[i * d + j] = s;
        }
}

/* ProductQuantizer::compute_inner_prod_table, faiss/ProductQuantizer.cpp:424-436
 * pqc layout [M][ksub][dsub] (ProductQuantizer.h get_centroids) */
void port_inner_prod_table(const float *q, const float *pqc, size_t M, size_t ksub, size_t dsub, float *table)
{
    for (size_t m = 0; m < M; m++)
        for (size_t j = 0; j < ksub; j++)
            table[m * ksub + j] = port_fvec_inner_product(q + m * dsub, pqc + (m * ksub + j) * dsub, dsub);
}

/* ------------------------------------------------------------- faiss heap */
/* faiss::maxheap_pop / maxheap_push, faiss/Heap.h:89-143 (CMax: cmp = a > b) */
static void heap_pop(size_t k, float *v, long *id)
{
    v--; id--;
    float val = v[k];
    size_t i = 1, i1, i2;
    while (1) {
        i1 = i << 1; i2 = i1 + 1;
        if (i1 > k) break;
        if (i2 == k + 1 || v[i1] > v[i2]) {
            if (val > v[i1]) break;
            v[i] = v[i1]; id[i] = id[i1]; i = i1;
        } else {
            if (val > v[i2]) break;
            v[i] = v[i2]; id[i] = id[i2]; i = i2;
        }
    }
    v[i] = v[k]; id[i] = id[k];
}
static void heap_push(size_t k, float *v, long *id, float val, long ids)
{
    v--; id--;
    size_t i = k, f;
    while (i > 1) {
        f = i >> 1;
        if (!(val > v[f])) break;
        v[i] = v[f]; id[i] = id[f]; i = f;
    }
    v[i] = val; id[i] = ids;
}
/* maxheap_heapify with k0 = 0, faiss/Heap.h:185-233 */
static void heap_init(size_t k, float *v, long *id)
{
    for (size_t i = 0; i < k; i++) { v[i] = FLT_MAX; id[i] = -1; }
}

/* IndexIVF_HNSW::pq_L2sqr, IndexIVF_HNSW.cpp:802-814 (sequential m order) */
static float pq_sum(const float *table, const uint8_t *code, size_t M, size_t ksub)
{
    float r = 0.f;
    for (size_t m = 0; m < M; m++) r += table[ksub * m + code[m]];
    return r;
}

/* ------------------------------------------------------------ coarse stage */
/* What the graft computes in place of hnswlib searchKnn (IndexIVF_HNSW.cpp:248-259):
 * the EXACT nprobe nearest centroids, ascending (dist, idx) like the
 * priority_queue<pair<float,idx_t>> unload order, with the reference's
 * fstdistfunc numerics.  centroids row-major [nc][d]. */
void port_coarse_exact(const float *q, const float *centroids, size_t nc, size_t d, size_t nprobe,
                       float *dists, uint32_t *idxs)
{
    size_t have = 0;
    for (size_t c = 0; c < nc; c++) {
        float dist = port_fvec_L2sqr16(q, centroids + c * d, d);
        /* insertion into ascending (dist, idx) array of size <= nprobe */
        size_t pos = have;
        while (pos > 0 && (dists[pos - 1] > dist)) pos--;
        if (pos >= nprobe) continue;
        size_t last = have < nprobe ? have : nprobe - 1;
        for (size_t t = last; t > pos; t--) { dists[t] = dists[t - 1]; idxs[t] = idxs[t - 1]; }
        dists[pos] = dist; idxs[pos] = (uint32_t)c;
        if (have < nprobe) have++;
    }
}

/* IndexIVF_HNSW::assign with k = 1 (IndexIVF_HNSW.cpp:68-72), exact version */
void port_assign_exact(size_t n, const float *x, const float *centroids, size_t nc, size_t d, uint32_t *labels)
{
    for (size_t i = 0; i < n; i++) {
        float best = 0.f; uint32_t bi = 0;
        for (size_t c = 0; c < nc; c++) {
            float dist = port_fvec_L2sqr16(x + i * d, centroids + c * d, d);
            if (c == 0 || dist < best) { best = dist; bi = (uint32_t)c; }
        }
        labels[i] = bi;
    }
}

/* --------------------------------------------------------- IVFADC search */
/* IndexIVF_HNSW::search2 (IndexIVF_HNSW.cpp:453-492) == the body of search
 * (:262-293) after the coarse stage.  query must already be OPQ-rotated by the
 * caller when do_opq (IndexIVF_HNSW.cpp:240).  Output is the faiss max-heap. */
void port_search_adc(size_t d, size_t M, size_t ksub, size_t nprobe, size_t max_codes, size_t k,
                     const float *query, const float *coarse_dists, const uint32_t *coarse_idxs,
                     const float *pqc, const float *norm_table, const float *centroid_norms,
                     const int64_t *list_off, const uint32_t *ids, const uint8_t *codes,
                     const uint8_t *norm_codes, float *distances, long *labels, size_t *ncode_out)
{
    const size_t dsub = d / M, code_size = M;
    float *table = (float *)malloc(M * ksub * sizeof(float));
    port_inner_prod_table(query, pqc, M, ksub, dsub, table);
    heap_init(k, distances, labels);
    size_t ncode = 0;
    for (size_t i = 0; i < nprobe; i++) {
        const uint32_t c = coarse_idxs[i];
        const size_t beg = (size_t)list_off[c], gs = (size_t)(list_off[c + 1] - list_off[c]);
        if (gs == 0) continue;
        const float term1 = coarse_dists[i] - centroid_norms[c];
        for (size_t j = 0; j < gs; j++) {
            const float term3 = 2 * pq_sum(table, codes + (beg + j) * code_size, M, ksub);
            const float dist = term1 + norm_table[norm_codes[beg + j]] - term3;
            if (dist < distances[0]) {
                heap_pop(k, distances, labels);
                heap_push(k, distances, labels, dist, (long)ids[beg + j]);
            }
        }
        ncode += gs;
        if (ncode >= max_codes) break;
    }
    if (ncode_out) *ncode_out = ncode;
    free(table);
}

/* ------------------------------------------------ Grouping (+pruning) search */
/* IndexIVF_HNSW_Grouping::search, IndexIVF_HNSW_Grouping.cpp:188-363, with the
 * quirks Q1(i)-(iv) of SURVEY.md 8(a) kept: qsd not advanced over empty lists,
 * zero qsd past the 2*max_codes horizon, `< EPS` meaning "not computed".
 * scratch_qcd must be an nc-sized array of zeros (the member
 * query_centroid_dists, IndexIVF_HNSW_Grouping.h:57); it is zeroed again on exit.
 * Optional debug outputs: threshold_out, scanned[nprobe*nsubc] (1 = sub-group scanned). */
void port_search_grouping(size_t d, size_t M, size_t ksub, size_t nsubc, size_t nprobe, size_t max_codes,
                          int do_pruning, size_t k, const float *query, const float *coarse_dists,
                          const uint32_t *coarse_idxs, const float *centroids, const float *pqc,
                          const float *norm_table, const float *centroid_norms, const int64_t *list_off,
                          const uint32_t *ids, const uint8_t *codes, const uint8_t *norm_codes,
                          const uint32_t *nn_idx, const uint32_t *sub_sizes, const float *alphas,
                          const float *icd, float *scratch_qcd, float *distances, long *labels,
                          float *threshold_out, uint8_t *scanned, size_t *ncode_out)
{
    const size_t dsub = d / M, code_size = M;
    uint32_t *used = (uint32_t *)malloc((nsubc * nprobe * 2 + nprobe) * sizeof(uint32_t));
    size_t nused = 0;
    float *qsd_buf = (float *)calloc(nsubc * nprobe, sizeof(float));
    float *table = (float *)malloc(M * ksub * sizeof(float));
    if (scanned) memset(scanned, 0, nprobe * nsubc);

    for (size_t i = 0; i < nprobe; i++) { /* :207-220 */
        scratch_qcd[coarse_idxs[i]] = coarse_dists[i];
        used[nused++] = coarse_idxs[i];
    }
    float threshold = 0.0f;
    if (do_pruning) { /* :223-262 */
        size_t ncode = 0, nsubgroups = 0;
        float *qsd = qsd_buf;
        for (size_t i = 0; i < nprobe; i++) {
            const uint32_t c = coarse_idxs[i];
            const size_t gs = (size_t)(list_off[c + 1] - list_off[c]);
            if (gs == 0) continue;
            const float alpha = alphas[c];
            const float term1 = (1 - alpha) * scratch_qcd[c];
            for (size_t s = 0; s < nsubc; s++) {
                if (sub_sizes[c * nsubc + s] == 0) continue;
                const uint32_t n = nn_idx[c * nsubc + s];
                if (scratch_qcd[n] < PORT_EPS) {
                    scratch_qcd[n] = port_fvec_L2sqr16(query, centroids + (size_t)n * d, d);
                    used[nused++] = n;
                }
                qsd[s] = term1 - alpha * ((1 - alpha) * icd[c * nsubc + s] - scratch_qcd[n]);
                threshold += qsd[s];
                nsubgroups++;
            }
            ncode += gs;
            qsd += nsubc;
            if (ncode >= 2 * max_codes) break;
        }
        threshold /= nsubgroups;
    }
    port_inner_prod_table(query, pqc, M, ksub, dsub, table);
    heap_init(k, distances, labels);

    size_t ncode = 0;
    const float *qsd = qsd_buf;
    for (size_t i = 0; i < nprobe; i++) { /* :283-353 */
        const uint32_t c = coarse_idxs[i];
        const size_t gs = (size_t)(list_off[c + 1] - list_off[c]);
        if (gs == 0) continue;
        const float alpha = alphas[c];
        const float term1 = (1 - alpha) * (scratch_qcd[c] - centroid_norms[c]);
        size_t pos = (size_t)list_off[c];
        for (size_t s = 0; s < nsubc; s++) {
            const size_t ss = sub_sizes[c * nsubc + s];
            if (ss == 0) continue;
            if (!do_pruning || qsd[s] < threshold) {
                const uint32_t n = nn_idx[c * nsubc + s];
                if (scratch_qcd[n] < PORT_EPS) {
                    scratch_qcd[n] = port_fvec_L2sqr16(query, centroids + (size_t)n * d, d);
                    used[nused++] = n;
                }
                const float term2 = alpha * (scratch_qcd[n] - centroid_norms[n]);
                for (size_t j = 0; j < ss; j++) {
                    const float term4 = 2 * pq_sum(table, codes + (pos + j) * code_size, M, ksub);
                    const float dist = term1 + term2 + norm_table[norm_codes[pos + j]] - term4;
                    if (dist < distances[0]) {
                        heap_pop(k, distances, labels);
                        heap_push(k, distances, labels, dist, (long)ids[pos + j]);
                    }
                }
                ncode += ss;
                if (scanned) scanned[i * nsubc + s] = 1;
            }
            pos += ss;
        }
        if (ncode >= max_codes) break;
        if (do_pruning) qsd += nsubc;
    }
    for (size_t u = 0; u < nused; u++) scratch_qcd[used[u]] = 0; /* :358-359 */
    if (threshold_out) *threshold_out = threshold;
    if (ncode_out) *ncode_out = ncode;
    free(table); free(qsd_buf); free(used);
}

/* -------------------------------------------------------------- indexing */
/* ProductQuantizer::compute_code(s), faiss/ProductQuantizer.cpp:311-336,385-407.
 * dsub < 16: direct ||x-c||^2 (fvec_L2sqr_ny); dsub >= 16: ||x||^2+||c||^2-2x.c
 * (pairwise_L2sqr, faiss/utils.cpp:1344-1388).  argmin strict <, first wins. */
void port_pq_encode(size_t n, size_t d, size_t M, size_t ksub, const float *pqc, const float *x, uint8_t *codes)
{
    const size_t dsub = d / M;
    for (size_t i = 0; i < n; i++)
        for (size_t m = 0; m < M; m++) {
            const float *xs = x + i * d + m * dsub;
            float mind = 1e20f; int best = -1;
            float xn = dsub >= 16 ? port_fvec_norm_L2sqr(xs, dsub) : 0.f;
            for (size_t j = 0; j < ksub; j++) {
                const float *c = pqc + (m * ksub + j) * dsub;
                float dis;
                if (dsub < 16) dis = port_faiss_L2sqr(xs, c, dsub);
                else {
                    float ip = 0.f;
                    for (size_t t = 0; t < dsub; t++) ip += xs[t] * c[t];
                    dis = (xn + port_fvec_norm_L2sqr(c, dsub)) + -2.0f * ip;
                }
                if (dis < mind) { mind = dis; best = (int)j; }
            }
            codes[i * M + m] = (uint8_t)best;
        }
}

/* ProductQuantizer::decode, faiss/ProductQuantizer.cpp:338-360 */
void port_pq_decode(size_t n, size_t d, size_t M, size_t ksub, const float *pqc, const uint8_t *codes, float *x)
{
    const size_t dsub = d / M;
    for (size_t i = 0; i < n; i++)
        for (size_t m = 0; m < M; m++)
            memcpy(x + i * d + m * dsub, pqc + (m * ksub + codes[i * M + m]) * dsub, dsub * sizeof(float));
}

/* Tail of add_batch / add_group shared by both (IndexIVF_HNSW.cpp:85-121,
 * IndexIVF_HNSW_Grouping.cpp:93-130): residual against `base` rows (centroid or
 * sub-centroid of each vector), optional OPQ, encode, decode, inverse OPQ,
 * reconstruct, norm, 1-D norm code.  Outputs codes[n*M], norm_codes[n],
 * optional recon_norms[n]. */
void port_encode_residuals(size_t n, size_t d, size_t M, size_t ksub, const float *pqc, const float *norm_table,
                           const float *opqA /* NULL = no opq */, const float *x, const float *base_rows,
                           uint8_t *codes, uint8_t *norm_codes, float *recon_norms)
{
    float *res = (float *)malloc(d * sizeof(float)), *tmp = (float *)malloc(d * sizeof(float));
    float *dec = (float *)malloc(d * sizeof(float));
    for (size_t i = 0; i < n; i++) {
        const float *b = base_rows + i * d;
        for (size_t j = 0; j < d; j++) res[j] = x[i * d + j] + -1.0f * b[j]; /* fvec_madd, faiss/utils.cpp:1866-1897 */
        if (opqA) { port_opq_apply(1, d, opqA, res, tmp); memcpy(res, tmp, d * sizeof(float)); }
        port_pq_encode(1, d, M, ksub, pqc, res, codes + i * M);
        port_pq_decode(1, d, M, ksub, pqc, codes + i * M, dec);
        if (opqA) { port_opq_apply_transpose(1, d, opqA, dec, tmp); memcpy(dec, tmp, d * sizeof(float)); }
        for (size_t j = 0; j < d; j++) dec[j] = dec[j] + 1.0f * b[j];
        float nr = port_fvec_norm_L2sqr(dec, d);
        if (recon_norms) recon_norms[i] = nr;
        /* norm_pq->compute_codes: d=1, M=1, dsub=1 -> direct (x-c)^2 argmin */
        float mind = 1e20f; int best = -1;
        for (size_t j = 0; j < ksub; j++) {
            float df = nr - norm_table[j]; float dis = df * df;
            if (dis < mind) { mind = dis; best = (int)j; }
        }
        norm_codes[i] = (uint8_t)best;
    }
    free(res); free(tmp); free(dec);
}

/* IndexIVF_HNSW::add_batch, IndexIVF_HNSW.cpp:75-138: encode against centroid
 * idx[i]; the scatter into lists (stable, input order) is done by the caller
 * with a stable counting sort (numpy argsort kind='stable'). */
void port_add_batch_encode(size_t n, size_t d, size_t M, size_t ksub, const float *pqc, const float *norm_table,
                           const float *opqA, const float *centroids, const float *x, const uint32_t *idx,
                           uint8_t *codes, uint8_t *norm_codes, float *recon_norms)
{
    float *rows = (float *)malloc(n * d * sizeof(float));
    for (size_t i = 0; i < n; i++) memcpy(rows + i * d, centroids + (size_t)idx[i] * d, d * sizeof(float));
    port_encode_residuals(n, d, M, ksub, pqc, norm_table, opqA, x, rows, codes, norm_codes, recon_norms);
    free(rows);
}

/* IndexIVF_HNSW_Grouping::compute_alpha, IndexIVF_HNSW_Grouping.cpp:691-728.
 * cv = centroid_vectors [nsubc][d] (n_s - c), cvn = ||cv_s||^2 as returned by
 * the HNSW (fstdistfunc numerics).  The std::priority_queue keeps the max of
 * (-dist, (num, den)) pairs, i.e. the smallest dist; pair ordering breaks dist
 * ties by larger numerator, then larger denominator. */
float port_compute_alpha(size_t d, size_t nsubc, const float *cv, const float *cvn, const float *centroid,
                         const float *points, size_t gs)
{
    float gnum = 0.f, gden = 0.f;
    float *pv = (float *)malloc(d * sizeof(float)), *sc = (float *)malloc(d * sizeof(float));
    for (size_t i = 0; i < gs; i++) {
        const float *p = points + i * d;
        for (size_t j = 0; j < d; j++) pv[j] = p[j] + -1.0f * centroid[j];
        int have = 0; float bnd = 0, bnum = 0, bden = 0; /* best = max of (-dist,num,den) */
        for (size_t s = 0; s < nsubc; s++) {
            const float *v = cv + s * d;
            float num = port_fvec_inner_product(v, pv, d);
            num = num > 0 ? num : 0.0f;
            const float den = cvn[s];
            const float a = num / den;
            for (size_t j = 0; j < d; j++) sc[j] = centroid[j] + a * v[j];
            const float nd = -port_fvec_L2sqr16(p, sc, d);
            int better = !have || nd > bnd || (nd == bnd && (num > bnum || (num == bnum && den > bden)));
            if (better) { have = 1; bnd = nd; bnum = num; bden = den; }
        }
        gnum += bnum; gden += bden;
    }
    free(pv); free(sc);
    return gden > 0 ? gnum / gden : 0.0f;
}

/* IndexIVF_HNSW_Grouping::add_group, IndexIVF_HNSW_Grouping.cpp:43-157, given
 * the group's neighbour list (nn[nsubc], nearest first, self excluded) and the
 * neighbour distances cvn[nsubc] (what the HNSW returned, :52-63).
 * Outputs, for the gs points in *sub-group order* (stable within a sub-group,
 * :133-156): out_ids, out_codes, out_norm_codes, sub_sizes[nsubc]; alpha_out;
 * sub_idx[gs] (in input order, for checks). */
void port_add_group(size_t d, size_t M, size_t ksub, size_t nsubc, const float *pqc, const float *norm_table,
                    const float *opqA, const float *centroids, uint32_t g, const uint32_t *nn, const float *cvn,
                    size_t gs, const float *x, const uint32_t *ids, float *alpha_out, uint32_t *sub_idx,
                    uint32_t *out_ids, uint8_t *out_codes, uint8_t *out_norm_codes, uint32_t *sub_sizes)
{
    memset(sub_sizes, 0, nsubc * sizeof(uint32_t));
    *alpha_out = 0.f;
    if (gs == 0) return; /* :64-65 (subgroup_sizes stays empty in the reference) */
    const float *centroid = centroids + (size_t)g * d;
    float *cv = (float *)malloc(nsubc * d * sizeof(float)), *subc = (float *)malloc(nsubc * d * sizeof(float));
    for (size_t s = 0; s < nsubc; s++)
        for (size_t j = 0; j < d; j++) cv[s * d + j] = centroids[(size_t)nn[s] * d + j] + -1.0f * centroid[j];
    const float alpha = port_compute_alpha(d, nsubc, cv, cvn, centroid, x, gs);
    *alpha_out = alpha;
    for (size_t s = 0; s < nsubc; s++)
        for (size_t j = 0; j < d; j++) subc[s * d + j] = centroid[j] + alpha * cv[s * d + j];
    /* compute_subcentroid_idxs, :673-689 */
    float *rows = (float *)malloc(gs * d * sizeof(float));
    for (size_t i = 0; i < gs; i++) {
        float mind = 0.f; long mi = -1;
        for (size_t s = 0; s < nsubc; s++) {
            float dist = port_fvec_L2sqr16(subc + s * d, x + i * d, d);
            if (mi == -1 || dist < mind) { mind = dist; mi = (long)s; }
        }
        sub_idx[i] = (uint32_t)mi;
        memcpy(rows + i * d, subc + (size_t)mi * d, d * sizeof(float));
    }
    uint8_t *codes = (uint8_t *)malloc(gs * M), *ncodes = (uint8_t *)malloc(gs);
    port_encode_residuals(gs, d, M, ksub, pqc, norm_table, opqA, x, rows, codes, ncodes, NULL);
    size_t w = 0;
    for (size_t s = 0; s < nsubc; s++)
        for (size_t i = 0; i < gs; i++)
            if (sub_idx[i] == s) {
                out_ids[w] = ids[i];
                memcpy(out_codes + w * M, codes + i * M, M);
                out_norm_codes[w] = ncodes[i];
                sub_sizes[s]++; w++;
            }
    free(cv); free(subc); free(rows); free(codes); free(ncodes);
}

/* Exact replacement of quantizer->searchKnn(centroid, nsubc+1) in add_group
 * (:47-63): the nsubc nearest OTHER centroids, nearest first, with
 * fstdistfunc distances.  (The reference drops the last-popped = nearest hit
 * assuming it is the centroid itself; we drop by id.) */
void port_centroid_neighbours(const float *centroids, size_t nc, size_t d, size_t nsubc, uint32_t g,
                              uint32_t *nn, float *cvn)
{
    float *dd = (float *)malloc((nsubc + 1) * sizeof(float));
    uint32_t *ii = (uint32_t *)malloc((nsubc + 1) * sizeof(uint32_t));
    size_t have = 0;
    for (size_t c = 0; c < nc; c++) {
        if (c == g) continue;
        float dist = port_fvec_L2sqr16(centroids + (size_t)g * d, centroids + c * d, d);
        size_t pos = have;
        while (pos > 0 && dd[pos - 1] > dist) pos--;
        if (pos >= nsubc) continue;
        size_t last = have < nsubc ? have : nsubc - 1;
        for (size_t t = last; t > pos; t--) { dd[t] = dd[t - 1]; ii[t] = ii[t - 1]; }
        dd[pos] = dist; ii[pos] = (uint32_t)c;
        if (have < nsubc) have++;
    }
    memcpy(nn, ii, nsubc * sizeof(uint32_t));
    memcpy(cvn, dd, nsubc * sizeof(float));
    free(dd); free(ii);
}

/* IndexIVF_HNSW::compute_centroid_norms, IndexIVF_HNSW.cpp:781-787 */
void port_centroid_norms(const float *centroids, size_t nc, size_t d, float *out)
{
    for (size_t c = 0; c < nc; c++) out[c] = port_fvec_norm_L2sqr(centroids + c * d, d);
}

/* IndexIVF_HNSW_Grouping::compute_inter_centroid_dists, IndexIVF_HNSW_Grouping.cpp:620-631 */
void port_inter_centroid_dists(const float *centroids, size_t nc, size_t d, size_t nsubc, const uint32_t *nn_idx,
                               float *icd)
{
    for (size_t c = 0; c < nc; c++)
        for (size_t s = 0; s < nsubc; s++)
            icd[c * nsubc + s] = port_fvec_L2sqr16(centroids + (size_t)nn_idx[c * nsubc + s] * d, centroids + c * d, d);
}
