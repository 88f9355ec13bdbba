// Indexing kernels: residual PQ encoding (K-B1), Grouping alpha / sub-centroid assignment (K-G2),
// inter-centroid distances (K-G3 tail) and the CSR list build (K-B2).
//
// Replaces the arithmetic of
//   IndexIVF_HNSW::add_batch             IndexIVF_HNSW.cpp:75-138
//   IndexIVF_HNSW_Grouping::add_group    IndexIVF_HNSW_Grouping.cpp:43-157
//   ::compute_alpha / ::compute_subcentroid_idxs   IndexIVF_HNSW_Grouping.cpp:673-728
//   ::compute_inter_centroid_dists       IndexIVF_HNSW_Grouping.cpp:620-631
//   faiss::ProductQuantizer::compute_codes / decode   faiss/ProductQuantizer.cpp:311-407
#include "common.cuh"
#include "kernels.h"

namespace ivfb200 {

// ---------------------------------------------------------------------------------------------
// out[i] = in[i] + sign * base_row(i)          (faiss::fvec_madd, faiss/utils.cpp:1866-1897)
// base_row(i) = centroid[idx[i]]                                   (IndexIVF_HNSW.cpp:817-831)
//             = c + alpha[g] * (centroid[nn_idx[g][sub[i]]] - c)   (IndexIVF_HNSW_Grouping.cpp:71-87,655-671)
// ---------------------------------------------------------------------------------------------
__global__ void base_combine_kernel(size_t n, int d, const float *__restrict__ in, const uint32_t *__restrict__ idx,
                                    const uint8_t *__restrict__ sub, const float *__restrict__ centroids,
                                    const uint32_t *__restrict__ nn_idx, const float *__restrict__ alphas,
                                    int nsubc, float sign, float *__restrict__ out)
{
    const size_t total = n * (size_t)d;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const size_t i = e / d;
        const int j = (int)(e - i * d);
        const uint32_t g = idx[i];
        float b = centroids[(size_t)g * d + j];
        if (sub) {
            const uint32_t nn = nn_idx[(size_t)g * nsubc + sub[i]];
            const float cv = __fadd_rn(centroids[(size_t)nn * d + j], -b);  // fvec_madd(nn, -1, c)
            b = __fadd_rn(b, __fmul_rn(alphas[g], cv));                      // fvec_madd(c, alpha, cv)
        }
        out[e] = __fadd_rn(in[e], sign * b);
    }
}

cudaError_t launch_residual(cudaStream_t st, uint64_t *launches, size_t n, int d, const float *x,
                            const uint32_t *idx, const uint8_t *sub, const float *centroids, const uint32_t *nn_idx,
                            const float *alphas, int nsubc, float sign, const float *addend, float *out)
{
    if (n == 0) return cudaSuccess;
    const float *in = addend ? addend : x;
    size_t blocks = (n * (size_t)d + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    base_combine_kernel<<<(unsigned)blocks, 256, 0, st>>>(n, d, in, idx, sub, centroids, nn_idx, alphas, nsubc, sign, out);
    ++*launches;
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// PQ encode: one thread per vector, the whole codebook in shared memory (all lanes of a warp look
// at the same (m, j) entry -> broadcast reads).  argmin is strict <, first wins
// (ProductQuantizer::compute_code, faiss/ProductQuantizer.cpp:311-336).  The distance follows the
// lane structure of faiss::fvec_L2sqr (faiss/utils.cpp:534-568).
// ---------------------------------------------------------------------------------------------
template <int DSUB>
static __device__ __forceinline__ float faiss_l2sqr_small(const float *x, const float *y)
{
    float l8[8];
#pragma unroll
    for (int l = 0; l < 8; l++) l8[l] = 0.f;
    int i = 0;
#pragma unroll
    for (; i + 8 <= DSUB; i += 8)
#pragma unroll
        for (int l = 0; l < 8; l++) { const float df = x[i + l] - y[i + l]; l8[l] = fmaf(df, df, l8[l]); }
    float l4[4];
#pragma unroll
    for (int l = 0; l < 4; l++) l4[l] = l8[l + 4] + l8[l];
    if (DSUB - i >= 4) {
#pragma unroll
        for (int l = 0; l < 4; l++) { const float df = x[i + l] - y[i + l]; l4[l] = fmaf(df, df, l4[l]); }
        i += 4;
    }
#pragma unroll
    for (int l = 0; l < 3; l++)
        if (i + l < DSUB) { const float df = x[i + l] - y[i + l]; l4[l] = fmaf(df, df, l4[l]); }
    return (l4[0] + l4[1]) + (l4[2] + l4[3]);
}

template <int DSUB>
__global__ void __launch_bounds__(256) pq_encode_kernel(size_t n, int M, const float *__restrict__ pqc,
                                                        const float *__restrict__ res, uint8_t *__restrict__ codes,
                                                        int cb_in_smem)
{
    extern __shared__ __align__(16) float scb[];  // [M][256][DSUB]
    const int d = M * DSUB;
    if (cb_in_smem) {
        for (int i = threadIdx.x; i < d * 256; i += blockDim.x) scb[i] = pqc[i];
        __syncthreads();
    }
    const float *cb = cb_in_smem ? scb : pqc;
    for (size_t v = (size_t)blockIdx.x * blockDim.x + threadIdx.x; v < n; v += (size_t)gridDim.x * blockDim.x) {
        for (int m = 0; m < M; m++) {
            float r[DSUB];
#pragma unroll
            for (int t = 0; t < DSUB; t++) r[t] = res[v * d + m * DSUB + t];
            float best = 1e20f;
            int bi = -1;
            const float *c = cb + (size_t)m * 256 * DSUB;
            for (int j = 0; j < 256; j++) {
                float y[DSUB];
#pragma unroll
                for (int t = 0; t < DSUB; t++) y[t] = c[j * DSUB + t];
                const float dis = faiss_l2sqr_small<DSUB>(r, y);
                if (dis < best) { best = dis; bi = j; }
            }
            codes[v * M + m] = (uint8_t)bi;
        }
    }
}

template <int DSUB>
static cudaError_t launch_pq_encode_t(cudaStream_t st, size_t n, int M, const float *pqc, const float *res,
                                      uint8_t *codes)
{
    const size_t cb_bytes = (size_t)M * 256 * DSUB * 4;
    const int in_smem = cb_bytes <= 200 * 1024;
    const size_t sm = in_smem ? cb_bytes : 0;
    cudaError_t e = cudaFuncSetAttribute(pq_encode_kernel<DSUB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)(sm > 48 * 1024 ? sm : 48 * 1024));
    if (e != cudaSuccess) return e;
    size_t blocks = (n + 255) / 256;
    if (blocks > 148) blocks = 148;
    pq_encode_kernel<DSUB><<<(unsigned)blocks, 256, sm, st>>>(n, M, pqc, res, codes, in_smem);
    return cudaGetLastError();
}

// any other sub-vector length: the same summation structure (8-lane blocks, a 4-lane block, a tail of <= 3) with a
// run-time dsub; slower (no register-resident sub-vector), bit-identical
static __device__ __forceinline__ float faiss_l2sqr_rt(const float *x, const float *y, int dsub)
{
    float l8[8];
#pragma unroll
    for (int l = 0; l < 8; l++) l8[l] = 0.f;
    int i = 0;
    for (; i + 8 <= dsub; i += 8)
#pragma unroll
        for (int l = 0; l < 8; l++) { const float df = x[i + l] - y[i + l]; l8[l] = fmaf(df, df, l8[l]); }
    float l4[4];
#pragma unroll
    for (int l = 0; l < 4; l++) l4[l] = l8[l + 4] + l8[l];
    if (dsub - i >= 4) {
#pragma unroll
        for (int l = 0; l < 4; l++) { const float df = x[i + l] - y[i + l]; l4[l] = fmaf(df, df, l4[l]); }
        i += 4;
    }
#pragma unroll
    for (int l = 0; l < 3; l++)
        if (i + l < dsub) { const float df = x[i + l] - y[i + l]; l4[l] = fmaf(df, df, l4[l]); }
    return (l4[0] + l4[1]) + (l4[2] + l4[3]);
}

__global__ void __launch_bounds__(256) pq_encode_generic_kernel(size_t n, int M, int dsub, const float *__restrict__ pqc,
                                                                const float *__restrict__ res,
                                                                uint8_t *__restrict__ codes)
{
    const int d = M * dsub;
    // one thread per (vector, sub-quantizer)
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < n * M; e += (size_t)gridDim.x * blockDim.x) {
        const size_t v = e / M;
        const int m = (int)(e - v * M);
        const float *r = res + v * d + (size_t)m * dsub;
        const float *c = pqc + (size_t)m * 256 * dsub;
        float best = 1e20f;
        int bi = -1;
        for (int j = 0; j < 256; j++) {
            const float dis = faiss_l2sqr_rt(r, c + (size_t)j * dsub, dsub);
            if (dis < best) { best = dis; bi = j; }
        }
        codes[e] = (uint8_t)bi;
    }
}

cudaError_t launch_pq_encode(cudaStream_t st, uint64_t *launches, size_t n, int d, int M, const float *pqc,
                             const float *residuals, uint8_t *codes)
{
    if (n == 0) return cudaSuccess;
    ++*launches;
    switch (d / M) {
        case 1: return launch_pq_encode_t<1>(st, n, M, pqc, residuals, codes);
        case 2: return launch_pq_encode_t<2>(st, n, M, pqc, residuals, codes);
        case 3: return launch_pq_encode_t<3>(st, n, M, pqc, residuals, codes);
        case 4: return launch_pq_encode_t<4>(st, n, M, pqc, residuals, codes);
        case 6: return launch_pq_encode_t<6>(st, n, M, pqc, residuals, codes);
        case 8: return launch_pq_encode_t<8>(st, n, M, pqc, residuals, codes);
        case 12: return launch_pq_encode_t<12>(st, n, M, pqc, residuals, codes);
        case 16: return launch_pq_encode_t<16>(st, n, M, pqc, residuals, codes);
        case 24: return launch_pq_encode_t<24>(st, n, M, pqc, residuals, codes);
        case 32: return launch_pq_encode_t<32>(st, n, M, pqc, residuals, codes);
    }
    {
        size_t blocks = (n * (size_t)M + 255) / 256;
        if (blocks > 148 * 8) blocks = 148 * 8;
        pq_encode_generic_kernel<<<(unsigned)blocks, 256, 0, st>>>(n, M, d / M, pqc, residuals, codes);
    }
    return cudaGetLastError();
}

// ProductQuantizer::decode (faiss/ProductQuantizer.cpp:338-360)
__global__ void pq_decode_kernel(size_t n, int d, int M, const float *__restrict__ pqc,
                                 const uint8_t *__restrict__ codes, float *__restrict__ out)
{
    const int dsub = d / M;
    const size_t total = n * (size_t)d;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const size_t i = e / d;
        const int j = (int)(e - i * d);
        const int m = j / dsub, t = j - m * dsub;
        out[e] = pqc[((size_t)m * 256 + codes[i * M + m]) * dsub + t];
    }
}

cudaError_t launch_pq_decode(cudaStream_t st, uint64_t *launches, size_t n, int d, int M, const float *pqc,
                             const uint8_t *codes, float *out)
{
    if (n == 0) return cudaSuccess;
    size_t blocks = (n * (size_t)d + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    pq_decode_kernel<<<(unsigned)blocks, 256, 0, st>>>(n, d, M, pqc, codes, out);
    ++*launches;
    return cudaGetLastError();
}

// norm of the reconstructed vector (faiss::fvec_norms_L2sqr, faiss/utils.cpp:706-713) and its 1-D code
// (norm_pq->compute_codes with d = M = dsub = 1: argmin_j (x - t_j)^2, strict <, first wins)
__global__ void __launch_bounds__(256) norm_encode_kernel(size_t n, int d, const float *__restrict__ recon,
                                                          const float *__restrict__ norm_table,
                                                          uint8_t *__restrict__ norm_codes, float *__restrict__ norms_out)
{
    __shared__ float st_[256];
    st_[threadIdx.x] = norm_table ? norm_table[threadIdx.x] : 0.f;
    __syncthreads();
    const int lane = threadIdx.x & 31, sub = lane & 3, b = lane & ~3;
    const size_t nthreads = (size_t)gridDim.x * blockDim.x;
    for (size_t t0 = (size_t)blockIdx.x * blockDim.x; t0 < n * 4; t0 += nthreads) {
        const size_t row = (t0 + threadIdx.x) >> 2;
        float acc = 0.f;
        if (row < n) {
            const float *r = recon + row * d;
            for (int i = sub; i < d; i += 4) acc = fmaf(r[i], r[i], acc);
        }
        const float s0 = __shfl_sync(IVF_FULL_MASK, acc, b), s1 = __shfl_sync(IVF_FULL_MASK, acc, b + 1);
        const float s2 = __shfl_sync(IVF_FULL_MASK, acc, b + 2), s3 = __shfl_sync(IVF_FULL_MASK, acc, b + 3);
        const float nr = (s0 + s1) + (s2 + s3);
        // the 4 lanes of a row split the 256 table entries; combine with (dis, j) lexicographic min
        float best = 1e20f;
        int bi = 256;
        for (int j = sub * 64; j < sub * 64 + 64; j++) {
            const float df = nr - st_[j];
            const float dis = df * df;
            if (dis < best) { best = dis; bi = j; }
        }
#pragma unroll
        for (int off = 1; off < 4; off <<= 1) {
            const float ob = __shfl_xor_sync(IVF_FULL_MASK, best, off);
            const int oi = __shfl_xor_sync(IVF_FULL_MASK, bi, off);
            if (ob < best || (ob == best && oi < bi)) { best = ob; bi = oi; }
        }
        if (row < n && sub == 0) {
            if (norm_codes) norm_codes[row] = (uint8_t)bi;
            if (norms_out) norms_out[row] = nr;
        }
    }
}

cudaError_t launch_norm_encode(cudaStream_t st, uint64_t *launches, size_t n, int d, const float *recon,
                               const float *norm_table, uint8_t *norm_codes, float *norms_out)
{
    if (n == 0) return cudaSuccess;
    size_t blocks = (n * 4 + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    norm_encode_kernel<<<(unsigned)blocks, 256, 0, st>>>(n, d, recon, norm_table, norm_codes, norms_out);
    ++*launches;
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// Grouping.  Points are addressed through `order` (sorted by group, stable), group g owns sorted
// positions [grp_off[g], grp_off[g+1]).  One CTA per group, lanes <-> sub-centroids.
// MODE 0: compute_alpha terms (IndexIVF_HNSW_Grouping.cpp:691-728): per point the (numerator,
//         denominator) of the sub-centroid direction that brings the projected point closest.
// MODE 1: compute_subcentroid_idxs (:673-689) with the final alpha.
// ---------------------------------------------------------------------------------------------
static __device__ __forceinline__ float faiss_inner_product_dev(const float *x, const float *y, int d, int ystride1)
{
    // faiss::fvec_inner_product (faiss/utils.cpp:500-532): 8 partial sums, folded to 4, tail, 2x hadd
    float l8[8];
#pragma unroll
    for (int l = 0; l < 8; l++) l8[l] = 0.f;
    int i = 0;
    for (; i + 8 <= d; i += 8)
#pragma unroll
        for (int l = 0; l < 8; l++) l8[l] = __fadd_rn(l8[l], __fmul_rn(x[(i + l) * ystride1], y[i + l]));
    float l4[4];
#pragma unroll
    for (int l = 0; l < 4; l++) l4[l] = l8[l + 4] + l8[l];
    if (d - i >= 4) {
#pragma unroll
        for (int l = 0; l < 4; l++) l4[l] = __fadd_rn(l4[l], __fmul_rn(x[(i + l) * ystride1], y[i + l]));
        i += 4;
    }
    for (int l = 0; i + l < d; l++) l4[l] = __fadd_rn(l4[l], __fmul_rn(x[(i + l) * ystride1], y[i + l]));
    return (l4[0] + l4[1]) + (l4[2] + l4[3]);
}

template <int MODE>
__global__ void __launch_bounds__(256) group_points_kernel(int d, int nsubc, const float *__restrict__ x,
                                                           const uint32_t *__restrict__ order,
                                                           const uint32_t *__restrict__ grp_off,
                                                           const float *__restrict__ centroids,
                                                           const uint32_t *__restrict__ nn_idx,
                                                           const float *__restrict__ icd,
                                                           const float *__restrict__ alphas, float *__restrict__ num_out,
                                                           float *__restrict__ den_out, uint8_t *__restrict__ sub_out)
{
    extern __shared__ __align__(16) float smf[];
    const int g = blockIdx.x;
    const uint32_t beg = grp_off[g], end = grp_off[g + 1];
    if (beg == end) return;
    const int ld = d + 1;
    float *cv = smf;                    // [nsubc][d+1]: centroid vectors (MODE 0) or sub-centroids (MODE 1)
    float *cen = cv + nsubc * ld;       // [d]
    float *cvn = cen + d;               // [nsubc]
    float *pbuf = cvn + nsubc;          // [8 warps][2][d]: point and point - centroid
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int j = tid; j < d; j += 256) cen[j] = centroids[(size_t)g * d + j];
    __syncthreads();
    const float alpha = MODE == 1 ? alphas[g] : 0.f;
    for (int e = tid; e < nsubc * d; e += 256) {
        const int s = e / d, j = e - s * d;
        const uint32_t nn = nn_idx[(size_t)g * nsubc + s];
        const float v = __fadd_rn(centroids[(size_t)nn * d + j], -cen[j]);           // n_s - c
        cv[s * ld + j] = MODE == 1 ? __fadd_rn(cen[j], __fmul_rn(alpha, v)) : v;      // c + alpha * cv
    }
    for (int s = tid; s < nsubc; s += 256) cvn[s] = icd[(size_t)g * nsubc + s];
    __syncthreads();
    float *pt = pbuf + warp * 2 * d, *pv = pt + d;
    const int n16 = (d >> 4) << 4;
    for (uint32_t r = beg + warp; r < end; r += 8) {
        const size_t src = order ? order[r] : r;
        __syncwarp();
        for (int j = lane; j < d; j += 32) {
            const float v = x[src * d + j];
            pt[j] = v;
            pv[j] = __fadd_rn(v, -cen[j]);
        }
        __syncwarp();
        // each lane handles sub-centroids s = lane, lane + 32, ...
        float b_nd = 0.f, b_num = 0.f, b_den = 0.f;
        int b_s = -1;
        for (int s = lane; s < nsubc; s += 32) {
            const float *row = cv + s * ld;
            if (MODE == 0) {
                float num = faiss_inner_product_dev(row, pv, d, 1);
                num = num > 0.f ? num : 0.f;
                const float den = cvn[s];
                const float a = num / den;
                float l8[8];
#pragma unroll
                for (int l = 0; l < 8; l++) l8[l] = 0.f;
                for (int i = 0; i < n16; i += 8)
#pragma unroll
                    for (int l = 0; l < 8; l++) {
                        const float sc = __fadd_rn(cen[i + l], __fmul_rn(a, row[i + l]));
                        const float df = pt[i + l] - sc;
                        l8[l] = __fadd_rn(l8[l], __fmul_rn(df, df));
                    }
                float dist = l8[0];
#pragma unroll
                for (int l = 1; l < 8; l++) dist += l8[l];
                const float nd = -dist;
                // std::priority_queue<pair<float, pair<float,float>>>::top(): lexicographic max
                const bool better = b_s < 0 || nd > b_nd || (nd == b_nd && (num > b_num || (num == b_num && den > b_den)));
                if (better) { b_s = s; b_nd = nd; b_num = num; b_den = den; }
            } else {
                float l8[8];
#pragma unroll
                for (int l = 0; l < 8; l++) l8[l] = 0.f;
                for (int i = 0; i < n16; i += 8)
#pragma unroll
                    for (int l = 0; l < 8; l++) {
                        const float df = row[i + l] - pt[i + l];
                        l8[l] = __fadd_rn(l8[l], __fmul_rn(df, df));
                    }
                float dist = l8[0];
#pragma unroll
                for (int l = 1; l < 8; l++) dist += l8[l];
                if (b_s < 0 || dist < b_nd) { b_s = s; b_nd = dist; }   // strict <, first (smallest s) wins
            }
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            const float o_nd = __shfl_xor_sync(IVF_FULL_MASK, b_nd, off);
            const float o_num = __shfl_xor_sync(IVF_FULL_MASK, b_num, off);
            const float o_den = __shfl_xor_sync(IVF_FULL_MASK, b_den, off);
            const int o_s = __shfl_xor_sync(IVF_FULL_MASK, b_s, off);
            bool take;
            if (o_s < 0) take = false;
            else if (b_s < 0) take = true;
            else if (MODE == 0)
                take = o_nd > b_nd || (o_nd == b_nd && (o_num > b_num || (o_num == b_num && o_den > b_den)));
            else take = o_nd < b_nd || (o_nd == b_nd && o_s < b_s);
            if (take) { b_nd = o_nd; b_num = o_num; b_den = o_den; b_s = o_s; }
        }
        if (lane == 0) {
            if (MODE == 0) { num_out[r] = b_num; den_out[r] = b_den; }
            else sub_out[src] = (uint8_t)b_s;
        }
    }
}

static size_t group_points_smem(int d, int nsubc)
{
    return (size_t)(nsubc * (d + 1) + d + nsubc + 8 * 2 * d) * sizeof(float);
}

cudaError_t launch_group_points(cudaStream_t st, uint64_t *launches, int mode, int nc, int d, int nsubc,
                                const float *x, const uint32_t *order, const uint32_t *grp_off,
                                const float *centroids, const uint32_t *nn_idx, const float *icd,
                                const float *alphas, float *num, float *den, uint8_t *sub)
{
    if (nc == 0) return cudaSuccess;
    const size_t sm = group_points_smem(d, nsubc);
    cudaError_t e;
    if (mode == 0) {
        e = cudaFuncSetAttribute(group_points_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        if (e != cudaSuccess) return e;
        group_points_kernel<0><<<nc, 256, sm, st>>>(d, nsubc, x, order, grp_off, centroids, nn_idx, icd, alphas, num,
                                                   den, sub);
    } else {
        e = cudaFuncSetAttribute(group_points_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        if (e != cudaSuccess) return e;
        group_points_kernel<1><<<nc, 256, sm, st>>>(d, nsubc, x, order, grp_off, centroids, nn_idx, icd, alphas, num,
                                                   den, sub);
    }
    ++*launches;
    return cudaGetLastError();
}

// alpha[g] = sum(num) / sum(den) over the group's points IN ORDER (sequential fp32 sums, :724-727)
__global__ void alpha_reduce_kernel(int nc, const uint32_t *__restrict__ grp_off, const float *__restrict__ num,
                                    const float *__restrict__ den, float *__restrict__ alphas)
{
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= nc) return;
    float sn = 0.f, sd = 0.f;
    for (uint32_t r = grp_off[g]; r < grp_off[g + 1]; r++) {
        sn = __fadd_rn(sn, num[r]);
        sd = __fadd_rn(sd, den[r]);
    }
    if (grp_off[g + 1] > grp_off[g]) alphas[g] = sd > 0.f ? sn / sd : 0.f;
}

cudaError_t launch_alpha_reduce(cudaStream_t st, uint64_t *launches, int nc, const uint32_t *grp_off,
                                const float *num, const float *den, float *alphas)
{
    if (nc == 0) return cudaSuccess;
    alpha_reduce_kernel<<<(nc + 127) / 128, 128, 0, st>>>(nc, grp_off, num, den, alphas);
    ++*launches;
    return cudaGetLastError();
}

// out[r][j] = L2sqr16(table[idx[r][j]], rows_x[r])   (8 lanes per distance)
__global__ void __launch_bounds__(256) gather_dists_kernel(size_t total, int per_row, int d,
                                                           const float *__restrict__ rows_x,
                                                           const uint32_t *__restrict__ idx,
                                                           const float *__restrict__ table, float *__restrict__ out)
{
    const int lane = threadIdx.x & 31, sub = lane & 7;
    const size_t octs = ((size_t)gridDim.x * blockDim.x) >> 3;
    const size_t rounds = (total + octs - 1) / octs;
    size_t e = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3;
    for (size_t it = 0; it < rounds; it++, e += octs) {
        const bool ok = e < total;
        const size_t r = ok ? e / per_row : 0;
        const uint32_t id = ok ? idx[e] : 0u;
        const float dist = l2sqr16_octet(table + (size_t)id * d, rows_x + r * d, d, sub);
        if (ok && sub == 0) out[e] = dist;
    }
}

cudaError_t launch_gather_dists(cudaStream_t st, uint64_t *launches, int nrow, int per_row, int d,
                                const float *rows_x, const uint32_t *idx, const float *table, float *out)
{
    const size_t total = (size_t)nrow * per_row;
    if (total == 0) return cudaSuccess;
    size_t blocks = (total * 8 + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    gather_dists_kernel<<<(unsigned)blocks, 256, 0, st>>>(total, per_row, d, rows_x, idx, table, out);
    ++*launches;
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// K-B2: CSR build.  `sorted_key[r]` / `order[r]` come from a stable radix sort of the records by
// key (= list id, or list*nsubc + sub-group for Grouping), so order of insertion survives inside a
// list exactly like the push_back loops of IndexIVF_HNSW.cpp:124-133 / IndexIVF_HNSW_Grouping.cpp:133-156.
// ---------------------------------------------------------------------------------------------
__global__ void hist_kernel(size_t n, const uint32_t *__restrict__ key, uint32_t *__restrict__ hist)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        atomicAdd(&hist[key[i]], 1u);
}

cudaError_t launch_hist(cudaStream_t st, uint64_t *launches, size_t n, const uint32_t *key, uint32_t *hist)
{
    if (n == 0) return cudaSuccess;
    size_t blocks = (n + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    hist_kernel<<<(unsigned)blocks, 256, 0, st>>>(n, key, hist);
    ++*launches;
    return cudaGetLastError();
}

// list_slot[c] = aligned exclusive prefix of list sizes; single CTA (nc <= a few million is fine)
// lists outside [keep_b, keep_e) get no slots (a sharded handle stores only the lists it owns); raw_off always counts
// every record, it is the position in the sorted staging order
__global__ void __launch_bounds__(1024) slot_scan_kernel(int nc, uint32_t align, const uint32_t *__restrict__ size,
                                                         uint32_t *__restrict__ slot, uint32_t *__restrict__ raw_off,
                                                         uint32_t *__restrict__ total_slots, int keep_b, int keep_e)
{
    __shared__ uint64_t s_a[1024], s_r[1024];
    __shared__ uint64_t carry_a, carry_r;
    if (threadIdx.x == 0) { carry_a = 0; carry_r = 0; }
    __syncthreads();
    for (int base = 0; base < nc; base += 1024) {
        const int i = base + threadIdx.x;
        const uint32_t sz = i < nc ? size[i] : 0u;
        const uint64_t al = (i >= keep_b && i < keep_e) ? ((uint64_t)sz + align - 1) / align * align : 0;
        s_a[threadIdx.x] = al;
        s_r[threadIdx.x] = sz;
        __syncthreads();
        for (int off = 1; off < 1024; off <<= 1) {
            uint64_t va = 0, vr = 0;
            if ((int)threadIdx.x >= off) { va = s_a[threadIdx.x - off]; vr = s_r[threadIdx.x - off]; }
            __syncthreads();
            s_a[threadIdx.x] += va;
            s_r[threadIdx.x] += vr;
            __syncthreads();
        }
        if (i < nc) {
            slot[i] = (uint32_t)(carry_a + s_a[threadIdx.x] - al);
            raw_off[i] = (uint32_t)(carry_r + s_r[threadIdx.x] - sz);
        }
        __syncthreads();
        if (threadIdx.x == 1023) { carry_a += s_a[1023]; carry_r += s_r[1023]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        total_slots[0] = (uint32_t)carry_a;
        total_slots[1] = (uint32_t)(carry_a >> 32);
        raw_off[nc] = (uint32_t)carry_r;
    }
}

cudaError_t launch_slot_scan(cudaStream_t st, uint64_t *launches, int nc, uint32_t align, const uint32_t *size,
                             uint32_t *slot, uint32_t *raw_off, uint32_t *total_slots, int keep_b, int keep_e)
{
    slot_scan_kernel<<<1, 1024, 0, st>>>(nc, align, size, slot, raw_off, total_slots, keep_b, keep_e);
    ++*launches;
    return cudaGetLastError();
}

// record at sorted position r -> slot list_slot[list] + (r - raw_off[list]); codes byte-permuted for the scan
__global__ void scatter_lists_kernel(size_t n, int M, int ml_mask, int key_div, const uint32_t *__restrict__ order,
                                     const uint32_t *__restrict__ sorted_key, const uint32_t *__restrict__ raw_off,
                                     const uint32_t *__restrict__ list_slot, const uint32_t *__restrict__ ids,
                                     const uint8_t *__restrict__ codes, const uint8_t *__restrict__ norm_codes,
                                     uint8_t *__restrict__ out_codes, uint8_t *__restrict__ out_norm,
                                     uint8_t *__restrict__ out_sub, uint32_t *__restrict__ out_ids, uint32_t keep_b,
                                     uint32_t keep_e)
{
    for (size_t r = (size_t)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (size_t)gridDim.x * blockDim.x) {
        const uint32_t key = sorted_key[r];
        const uint32_t list = key / key_div, sub = key - list * key_div;
        if (list < keep_b || list >= keep_e) continue;   // not resident on this shard
        const size_t src = order ? order[r] : r;
        const uint32_t slot = list_slot[list] + (uint32_t)(r - raw_off[list]);
        const uint32_t x = slot & ml_mask;
        for (int t = 0; t < M; t++) out_codes[(size_t)slot * M + t] = codes[src * M + (t ^ x)];
        out_norm[slot] = norm_codes[src];
        if (out_sub) out_sub[slot] = (uint8_t)sub;
        out_ids[slot] = ids[src];
    }
}

cudaError_t launch_scatter_lists2(cudaStream_t st, uint64_t *launches, size_t n, int M, int key_div,
                                  const uint32_t *order, const uint32_t *sorted_key, const uint32_t *raw_off,
                                  const uint32_t *list_slot, const uint32_t *ids, const uint8_t *codes,
                                  const uint8_t *norm_codes, uint8_t *out_codes, uint8_t *out_norm, uint8_t *out_sub,
                                  uint32_t *out_ids, uint32_t keep_b, uint32_t keep_e)
{
    if (n == 0) return cudaSuccess;
    const int ml = M < 32 ? M : 32;
    size_t blocks = (n + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    scatter_lists_kernel<<<(unsigned)blocks, 256, 0, st>>>(n, M, ml - 1, key_div, order, sorted_key, raw_off, list_slot,
                                                        ids, codes, norm_codes, out_codes, out_norm, out_sub, out_ids,
                                                        keep_b, keep_e);
    ++*launches;
    return cudaGetLastError();
}

// inverse of the scatter for export: slot layout -> row-major records in list order
__global__ void gather_lists_kernel(int nc, int M, int ml_mask, const uint32_t *__restrict__ list_size,
                                    const uint32_t *__restrict__ list_slot, const uint32_t *__restrict__ raw_off,
                                    const uint8_t *__restrict__ codes, const uint8_t *__restrict__ norm_codes,
                                    const uint32_t *__restrict__ ids, uint8_t *__restrict__ out_codes,
                                    uint8_t *__restrict__ out_norm, uint32_t *__restrict__ out_ids)
{
    const int c = blockIdx.x;
    const uint32_t n = list_size[c], s0 = list_slot[c], o0 = raw_off[c];
    for (uint32_t j = threadIdx.x; j < n; j += blockDim.x) {
        const uint32_t slot = s0 + j, x = slot & ml_mask;
        for (int t = 0; t < M; t++) out_codes[(size_t)(o0 + j) * M + (t ^ x)] = codes[(size_t)slot * M + t];
        out_norm[o0 + j] = norm_codes[slot];
        out_ids[o0 + j] = ids[slot];
    }
}

cudaError_t launch_gather_lists(cudaStream_t st, uint64_t *launches, int nc, int M, const uint32_t *list_size,
                                const uint32_t *list_slot, const uint32_t *raw_off, const uint8_t *codes,
                                const uint8_t *norm_codes, const uint32_t *ids, uint8_t *out_codes, uint8_t *out_norm,
                                uint32_t *out_ids)
{
    if (nc == 0) return cudaSuccess;
    const int ml = M < 32 ? M : 32;
    gather_lists_kernel<<<nc, 128, 0, st>>>(nc, M, ml - 1, list_size, list_slot, raw_off, codes, norm_codes, ids,
                                            out_codes, out_norm, out_ids);
    ++*launches;
    return cudaGetLastError();
}

}  // namespace ivfb200
