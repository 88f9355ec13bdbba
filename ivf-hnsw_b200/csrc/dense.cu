// Small dense helpers: OPQ rotation (K-R1), row norms (K-N), u8 -> f32 expansion.
#include "common.cuh"
#include "kernels.h"

namespace ivfb200 {

// LinearTransform::apply_noalloc / transform_transpose (faiss/VectorTransform.cpp:106-158) for a
// d x d matrix without bias.  ROT_ROWS rows per CTA staged in shared memory; thread o owns output column o and
// adds a_oj * x_j for j = 0, 1, ... in order (one fmaf chain per output, as before).  The matrix goes through
// shared memory in 128 x 32 tiles loaded along its contiguous dimension: reading A[o][j] straight from global,
// one row per thread, touched 32 lines per warp load and left the kernel L1-bound (100 us for 10k x 128).
#define ROT_ROWS 16
#define ROT_JC 32
template <bool TRANSPOSE>
__global__ void __launch_bounds__(128) rotate_kernel(size_t n, int d, const float *__restrict__ A,
                                                     const float *__restrict__ x, float *__restrict__ out)
{
    extern __shared__ __align__(16) float sx[];  // [ROT_ROWS][d] then the tile [128][ROT_JC + 1]
    float *sa = sx + (size_t)ROT_ROWS * d;
    const int tid = threadIdx.x;
    const size_t row0 = (size_t)blockIdx.x * ROT_ROWS;
    const int rows = (int)((n - row0) < ROT_ROWS ? (n - row0) : ROT_ROWS);
    for (int i = tid; i < ROT_ROWS * d; i += 128) sx[i] = i < rows * d ? x[row0 * d + i] : 0.f;
    for (int ob = 0; ob < d; ob += 128) {
        const int o = ob + tid;
        float acc[ROT_ROWS];
#pragma unroll
        for (int r = 0; r < ROT_ROWS; r++) acc[r] = 0.f;
        for (int j0 = 0; j0 < d; j0 += ROT_JC) {
            __syncthreads();   // previous tile consumed (and sx written, first time round)
            if (TRANSPOSE) {
                // a(o, j) = A[j][o]: contiguous along o
                for (int it = 0; it < ROT_JC; it++) {
                    const int j = j0 + it;
                    sa[tid * (ROT_JC + 1) + it] = (j < d && o < d) ? __ldg(A + (size_t)j * d + o) : 0.f;
                }
            } else {
                // a(o, j) = A[o][j]: contiguous along j
                for (int it = 0; it < 32; it++) {
                    const int oo = it * 4 + (tid >> 5), jj = tid & 31;
                    sa[oo * (ROT_JC + 1) + jj] = (ob + oo < d && j0 + jj < d) ? __ldg(A + (size_t)(ob + oo) * d + j0 + jj) : 0.f;
                }
            }
            __syncthreads();
            const int jn = min(ROT_JC, d - j0);
            for (int jj = 0; jj < jn; jj++) {
                const float a = sa[tid * (ROT_JC + 1) + jj];
#pragma unroll
                for (int r = 0; r < ROT_ROWS; r++) acc[r] = fmaf(a, sx[r * d + j0 + jj], acc[r]);
            }
        }
        if (o < d)
            for (int r = 0; r < rows; r++) out[(row0 + r) * d + o] = acc[r];
    }
}

cudaError_t launch_rotate(cudaStream_t st, uint64_t *launches, size_t n, int d, const float *A, bool transpose,
                          const float *x, float *out)
{
    if (n == 0) return cudaSuccess;
    const unsigned grid = (unsigned)((n + ROT_ROWS - 1) / ROT_ROWS);
    const size_t sm = ((size_t)ROT_ROWS * d + 128 * (ROT_JC + 1)) * sizeof(float);
    cudaError_t e;
    if (transpose) {
        e = cudaFuncSetAttribute(rotate_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        if (e != cudaSuccess) return e;
        rotate_kernel<true><<<grid, 128, sm, st>>>(n, d, A, x, out);
    } else {
        e = cudaFuncSetAttribute(rotate_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        if (e != cudaSuccess) return e;
        rotate_kernel<false><<<grid, 128, sm, st>>>(n, d, A, x, out);
    }
    ++*launches;
    return cudaGetLastError();
}

// faiss::fvec_norm_L2sqr (faiss/utils.cpp:632-650): 4 partial sums, (s0+s1)+(s2+s3).
// Four consecutive lanes per row.
__global__ void row_norms_kernel(size_t n, int d, int dk, const float *__restrict__ x, float *__restrict__ out)
{
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t row = t >> 2;
    const int sub = (int)(t & 3);
    float acc = 0.f;
    if (row < n) {
        const float *r = x + row * d;
        for (int i = sub; i < dk; i += 4) acc = fmaf(r[i], r[i], acc);
    }
    const int lane = threadIdx.x & 31, b = lane & ~3;
    const float s0 = __shfl_sync(IVF_FULL_MASK, acc, b), s1 = __shfl_sync(IVF_FULL_MASK, acc, b + 1);
    const float s2 = __shfl_sync(IVF_FULL_MASK, acc, b + 2), s3 = __shfl_sync(IVF_FULL_MASK, acc, b + 3);
    if (row < n && sub == 0) out[row] = (s0 + s1) + (s2 + s3);
}

cudaError_t launch_row_norms(cudaStream_t st, uint64_t *launches, size_t n, int d, const float *x, float *out, int dk)
{
    if (n == 0) return cudaSuccess;
    const size_t threads = n * 4;
    row_norms_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(n, d, dk < 0 ? d : dk, x, out);
    ++*launches;
    return cudaGetLastError();
}

__global__ void u8_to_f32_kernel(size_t n4, const uchar4 *__restrict__ x, float4 *__restrict__ out)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        const uchar4 v = x[i];
        out[i] = make_float4((float)v.x, (float)v.y, (float)v.z, (float)v.w);
    }
}

cudaError_t launch_u8_to_f32(cudaStream_t st, uint64_t *launches, size_t n, const uint8_t *x, float *out)
{
    if (n == 0) return cudaSuccess;
    const size_t n4 = n / 4;  // callers pass n = rows * d with d % 4 == 0
    size_t blocks = (n4 + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    u8_to_f32_kernel<<<(unsigned)blocks, 256, 0, st>>>(n4, reinterpret_cast<const uchar4 *>(x),
                                                     reinterpret_cast<float4 *>(out));
    ++*launches;
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// K-RR: exact re-rank of a query's candidates against their RAW vectors -- the second half of
// IndexIVF_HNSW_Grouping::searchDisk (IndexIVF_HNSW_Grouping.cpp:365-395): d(q, x) with
// ivfhnsw::fvec_L2sqr numerics (utils.cpp:22-51: floor(d/16)*16 dims, 8 partial sums), then ascending
// (distance, label).  One CTA per query, 8 lanes per candidate; candidates with label < 0 (padding of a
// short result list) sort last.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) rerank_kernel(int ncand, int k, int d, const float *__restrict__ Q,
                                                     const float *__restrict__ cand,
                                                     const int64_t *__restrict__ cand_labels,
                                                     float *__restrict__ out_d, int64_t *__restrict__ out_l)
{
    extern __shared__ __align__(16) float rr_sm[];
    float *sq = rr_sm;           // [d]
    float *sd = rr_sm + d;       // [ncand]
    const int q = blockIdx.x, tid = threadIdx.x, lane = tid & 31, sub = lane & 7;
    for (int i = tid; i < d; i += 256) sq[i] = Q[(size_t)q * d + i];
    __syncthreads();
    const int oct = tid >> 3;    // 32 octets per CTA
    for (int c0 = 0; c0 < ncand; c0 += 32) {
        const int c = c0 + oct;
        const int cc = c < ncand ? c : ncand - 1;
        const float dist = l2sqr16_octet(sq, cand + ((size_t)q * ncand + cc) * d, d, sub);
        if (c < ncand && sub == 0) sd[c] = cand_labels[(size_t)q * ncand + c] < 0 ? FLT_MAX : dist;
    }
    __syncthreads();
    for (int e = tid; e < ncand; e += 256) {
        const float ed = sd[e];
        const int64_t el = cand_labels[(size_t)q * ncand + e];
        int rank = 0;
        for (int o = 0; o < ncand; o++) {
            const float od = sd[o];
            const int64_t ol = cand_labels[(size_t)q * ncand + o];
            rank += (od < ed || (od == ed && (ol < el || (ol == el && o < e)))) ? 1 : 0;
        }
        if (rank < k) {
            out_d[(size_t)q * k + rank] = ed;
            out_l[(size_t)q * k + rank] = el < 0 ? -1 : el;
        }
    }
}

cudaError_t launch_rerank(cudaStream_t st, uint64_t *launches, int nq, int ncand, int k, int d, const float *queries,
                          const float *cand, const int64_t *cand_labels, float *out_d, int64_t *out_l)
{
    if (nq == 0) return cudaSuccess;
    const size_t sm = (size_t)(d + ncand) * sizeof(float);
    cudaError_t e = cudaFuncSetAttribute(rerank_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    if (e != cudaSuccess) return e;
    rerank_kernel<<<nq, 256, sm, st>>>(ncand, k, d, queries, cand, cand_labels, out_d, out_l);
    ++*launches;
    return cudaGetLastError();
}

}  // namespace ivfb200
