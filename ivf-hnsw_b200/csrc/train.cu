// train_pq on the GPU:
//   IndexIVF_HNSW::train_pq            IndexIVF_HNSW.cpp:536-593
//   IndexIVF_HNSW_Grouping::train_pq   IndexIVF_HNSW_Grouping.cpp:486-618
//   faiss::ProductQuantizer::train     faiss/ProductQuantizer.cpp:236-308   (M independent 256-way k-means)
//   faiss::Clustering::train           faiss/Clustering.cpp:67-230          (25 Lloyd iterations)
//   faiss::km_update_centroids         faiss/utils.cpp:1401-1484            (void-cluster splitting)
//   faiss::OPQMatrix::train            faiss/VectorTransform.cpp:587-744    (alternating PQ / Procrustes)
// The assignment step of every Lloyd iteration is the PQ encode kernel (exact nearest centroid), the
// accumulation is a double-precision atomic histogram; the tiny centroid update (void-cluster
// splitting with faiss' glibc random_r sequence) and the d x d SVD of OPQ run on the host.
#include <stdlib.h>

#include <cmath>
#include <cstring>

#include "common.cuh"
#include "handle.h"

using namespace ivfb200;

#define CU(call)                                                                                         \
    do {                                                                                                 \
        cudaError_t e__ = (call);                                                                        \
        if (e__ != cudaSuccess)                                                                          \
            return h->fail(IVFHNSW_B200_ECUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
    } while (0)
#define LAUNCHES (&h->stats.kernel_launches)

namespace {

// faiss::RandomGenerator on linux (faiss/utils.cpp:139-159) and rand_perm (:311-321)
struct FaissRng {
    char state[8];
    struct random_data rd;
    explicit FaissRng(long seed)
    {
        memset(&rd, 0, sizeof rd);
        initstate_r((unsigned)seed, state, sizeof state, &rd);
    }
    int rand_int()
    {
        int32_t a;
        random_r(&rd, &a);
        return a;
    }
    int rand_int(int max) { return rand_int() % max; }
    float rand_float() { return rand_int() / float(1L << 31); }
};

void faiss_rand_perm(std::vector<int> &perm, size_t n, long seed)
{
    perm.resize(n);
    for (size_t i = 0; i < n; i++) perm[i] = (int)i;
    FaissRng rng(seed);
    for (size_t i = 0; i + 1 < n; i++) {
        const int i2 = (int)i + rng.rand_int((int)(n - i));
        std::swap(perm[i], perm[i2]);
    }
}

__global__ void gather_rows_kernel(size_t n, int d, const float *__restrict__ x, const int *__restrict__ idx,
                                   float *__restrict__ out)
{
    const size_t total = n * (size_t)d;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const size_t i = e / d;
        out[e] = x[(size_t)idx[i] * d + (e - i * d)];
    }
}

// sums[m][code][t] += x[v][m*dsub + t]; counts[m][code] += 1
__global__ void kmeans_accum_kernel(size_t n, int d, int M, const float *__restrict__ x,
                                    const uint8_t *__restrict__ codes, double *__restrict__ sums,
                                    unsigned int *__restrict__ counts)
{
    const int dsub = d / M;
    const size_t total = n * (size_t)d;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const size_t v = e / d;
        const int j = (int)(e - v * d), m = j / dsub, t = j - m * dsub;
        const int c = codes[v * M + m];
        atomicAdd(&sums[((size_t)m * 256 + c) * dsub + t], (double)x[e]);
        if (t == 0) atomicAdd(&counts[m * 256 + c], 1u);
    }
}

// coarse k-means: sums[label][j] += x[v][j] (double), counts[label] += 1, obj += d(x, centroid)
__global__ void coarse_accum_kernel(size_t n, int d, const float *__restrict__ x, const uint32_t *__restrict__ lab,
                                    const float *__restrict__ dist, double *__restrict__ sums,
                                    unsigned int *__restrict__ counts, double *__restrict__ obj)
{
    const size_t total = n * (size_t)d;
    double o = 0;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const size_t v = e / d;
        const int j = (int)(e - v * d);
        const uint32_t c = lab[v];
        atomicAdd(&sums[(size_t)c * d + j], (double)x[e]);
        if (j == 0) { atomicAdd(&counts[c], 1u); o += (double)dist[v]; }
    }
    for (int off = 16; off > 0; off >>= 1) o += __shfl_xor_sync(0xffffffffu, o, off);
    if ((threadIdx.x & 31) == 0 && o != 0) atomicAdd(obj, o);
}
// centroid <- sum / count where the cluster is not void (void ones are re-seeded by the split list)
__global__ void coarse_mean_kernel(size_t k, int d, const double *__restrict__ sums, const unsigned int *__restrict__ counts,
                                   float *__restrict__ cent)
{
    const size_t total = k * (size_t)d;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const unsigned int n = counts[e / d];
        cent[e] = n ? (float)(sums[e] / (double)n) : 0.f;
    }
}
// km_update_centroids' void-cluster loop (faiss/utils.cpp:1450-1478) applied IN ORDER: pair p copies centroid cj over
// ci and perturbs both symmetrically; a cluster may be split more than once, hence one CTA walking the list
__global__ void coarse_split_kernel(int npairs, int d, const uint32_t *__restrict__ pairs, float *__restrict__ cent)
{
    const float eps = 1 / 1024.f;
    for (int p = 0; p < npairs; p++) {
        const size_t ci = pairs[2 * p], cj = pairs[2 * p + 1];
        for (int j = threadIdx.x; j < d; j += blockDim.x) {
            const float v = cent[cj * d + j];
            cent[ci * d + j] = v * ((j % 2 == 0) ? 1 + eps : 1 - eps);
            cent[cj * d + j] = v * ((j % 2 == 0) ? 1 - eps : 1 + eps);
        }
        __syncthreads();
    }
}

__global__ void center_kernel(size_t n, int d, const double *__restrict__ colsum, float *__restrict__ x)
{
    const size_t total = n * (size_t)d;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x)
        x[e] -= (float)(colsum[e % d] / (double)n);
}
__global__ void colsum_kernel(size_t n, int d, const float *__restrict__ x, double *__restrict__ colsum)
{
    // one block per column chunk; simple and only used by OPQ training
    const int j = blockIdx.x;
    double acc = 0;
    for (size_t i = threadIdx.x; i < n; i += blockDim.x) acc += x[i * d + j];
    for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
    if ((threadIdx.x & 31) == 0) atomicAdd(&colsum[j], acc);
}
// C[i][j] = sum_n Y[n][i] * X[n][j]   (d x d, double accumulation), 16x16 tile per block
__global__ void __launch_bounds__(256) cross_cov_kernel(size_t n, int d, const float *__restrict__ Y,
                                                        const float *__restrict__ X, double *__restrict__ Cm)
{
    __shared__ float sy[32][17], sx[32][17];
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const int i0 = blockIdx.y * 16, j0 = blockIdx.x * 16;
    double acc = 0;
    for (size_t r0 = 0; r0 < n; r0 += 32) {
        for (int t = threadIdx.x; t < 32 * 16; t += 256) {
            const int r = t >> 4, c = t & 15;
            const size_t row = r0 + r;
            sy[r][c] = (row < n && i0 + c < d) ? Y[row * d + i0 + c] : 0.f;
            sx[r][c] = (row < n && j0 + c < d) ? X[row * d + j0 + c] : 0.f;
        }
        __syncthreads();
        float part = 0.f;
#pragma unroll
        for (int r = 0; r < 32; r++) part = fmaf(sy[r][ty], sx[r][tx], part);
        acc += part;
        __syncthreads();
    }
    if (i0 + ty < d && j0 + tx < d) Cm[(size_t)(i0 + ty) * d + j0 + tx] = acc;
}

unsigned grid_for(size_t n, int per = 256, unsigned cap = 148 * 16)
{
    size_t b = (n + per - 1) / per;
    if (b < 1) b = 1;
    return (unsigned)std::min<size_t>(b, cap);
}

// km_update_centroids (faiss/utils.cpp:1401-1484) for one sub-quantizer, on the host
int km_update_host(int k, int dsub, size_t n, const double *sums, const unsigned *counts, float *cent)
{
    std::vector<size_t> hassign(k);
    for (int c = 0; c < k; c++) {
        hassign[c] = counts[c];
        if (counts[c])
            for (int t = 0; t < dsub; t++) cent[c * dsub + t] = (float)(sums[c * dsub + t] / (double)counts[c]);
        else
            for (int t = 0; t < dsub; t++) cent[c * dsub + t] = 0.f;
    }
    int nsplit = 0;
    FaissRng rng(1234);
    const float eps = 1 / 1024.f;
    for (int ci = 0; ci < k; ci++) {
        if (hassign[ci] != 0) continue;
        int cj;
        for (cj = 0; true; cj = (cj + 1) % k) {
            const float p = (hassign[cj] - 1.0f) / (float)(n - k);
            if (rng.rand_float() < p) break;
        }
        memcpy(cent + ci * dsub, cent + cj * dsub, sizeof(float) * dsub);
        for (int t = 0; t < dsub; t++) {
            if (t % 2 == 0) { cent[ci * dsub + t] *= 1 + eps; cent[cj * dsub + t] *= 1 - eps; }
            else { cent[ci * dsub + t] *= 1 - eps; cent[cj * dsub + t] *= 1 + eps; }
        }
        hassign[ci] = hassign[cj] / 2;
        hassign[cj] -= hassign[ci];
        nsplit++;
    }
    return nsplit;
}

// ProductQuantizer::train for M sub-spaces at once.  x_dev [n][d] device; cent_io host [M][256][dsub]
// (hot_start: use cent_io as the initial centroids, faiss Train_hot_start)
int pq_kmeans(ivfhnsw_b200_t *h, size_t n, int d, int M, const float *x_dev, int niter, size_t max_per_centroid,
              bool hot_start, std::vector<float> &cent_io, DevBuf &cent_dev)
{
    const int k = 256, dsub = d / M;
    cudaStream_t st = h->stream;
    if (n < (size_t)k)
        return h->fail(IVFHNSW_B200_EINVAL, "Number of training points (%zu) should be at least as large as number of clusters (256)", n);
    // Clustering::train sub-sampling (faiss/Clustering.cpp:84-95), same permutation for every slice
    const float *xt = x_dev;
    size_t nx = n;
    DevBuf sub, idxbuf;
    if (n > (size_t)k * max_per_centroid) {
        std::vector<int> perm;
        faiss_rand_perm(perm, n, 1234);
        nx = (size_t)k * max_per_centroid;
        CU(idxbuf.ensure(nx * 4));
        CU(sub.ensure(nx * (size_t)d * 4));
        CU(cudaMemcpyAsync(idxbuf.p, perm.data(), nx * 4, cudaMemcpyHostToDevice, st));
        gather_rows_kernel<<<grid_for(nx * d), 256, 0, st>>>(nx, d, x_dev, idxbuf.as<int>(), sub.as<float>());
        ++*LAUNCHES;
        CU(cudaStreamSynchronize(st));
        xt = sub.as<float>();
    }
    CU(cent_dev.ensure((size_t)d * k * 4));
    if (!hot_start) {
        // centroids <- x[perm[i]], i < k  (faiss/Clustering.cpp:158-162, seed + 1)
        std::vector<int> perm;
        faiss_rand_perm(perm, nx, 1234 + 1);
        DevBuf rows, ridx;
        CU(ridx.ensure(k * 4));
        CU(rows.ensure((size_t)k * d * 4));
        CU(cudaMemcpyAsync(ridx.p, perm.data(), k * 4, cudaMemcpyHostToDevice, st));
        gather_rows_kernel<<<grid_for((size_t)k * d), 256, 0, st>>>(k, d, xt, ridx.as<int>(), rows.as<float>());
        ++*LAUNCHES;
        std::vector<float> init((size_t)k * d);
        CU(cudaMemcpyAsync(init.data(), rows.p, (size_t)k * d * 4, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        cent_io.resize((size_t)d * k);
        for (int m = 0; m < M; m++)
            for (int c = 0; c < k; c++)
                memcpy(&cent_io[((size_t)m * k + c) * dsub], &init[(size_t)c * d + m * dsub], dsub * sizeof(float));
        rows.release();
        ridx.release();
    }
    DevBuf codes, sums, counts;
    CU(codes.ensure(nx * M));
    CU(sums.ensure((size_t)d * k * 8));
    CU(counts.ensure((size_t)M * k * 4));
    std::vector<double> hs((size_t)d * k);
    std::vector<unsigned> hc((size_t)M * k);
    for (int it = 0; it < niter; it++) {
        CU(cudaMemcpyAsync(cent_dev.p, cent_io.data(), (size_t)d * k * 4, cudaMemcpyHostToDevice, st));
        CU(launch_pq_encode(st, LAUNCHES, nx, d, M, cent_dev.as<float>(), xt, codes.as<uint8_t>()));
        CU(cudaMemsetAsync(sums.p, 0, (size_t)d * k * 8, st));
        CU(cudaMemsetAsync(counts.p, 0, (size_t)M * k * 4, st));
        kmeans_accum_kernel<<<grid_for(nx * d), 256, 0, st>>>(nx, d, M, xt, codes.as<uint8_t>(), sums.as<double>(),
                                                             counts.as<unsigned>());
        ++*LAUNCHES;
        CU(cudaMemcpyAsync(hs.data(), sums.p, (size_t)d * k * 8, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(hc.data(), counts.p, (size_t)M * k * 4, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        for (int m = 0; m < M; m++)
            km_update_host(k, dsub, nx, &hs[(size_t)m * k * dsub], &hc[(size_t)m * k], &cent_io[(size_t)m * k * dsub]);
    }
    CU(cudaMemcpyAsync(cent_dev.p, cent_io.data(), (size_t)d * k * 4, cudaMemcpyHostToDevice, st));
    CU(cudaStreamSynchronize(st));
    codes.release(); sums.release(); counts.release(); sub.release(); idxbuf.release();
    return 0;
}

// one-sided Jacobi SVD of a d x d matrix (row-major, double): Min = U diag(s) V^T.  Returns U and V.
void jacobi_svd(int d, std::vector<double> A, std::vector<double> &U, std::vector<double> &V)
{
    // work on columns of A (A <- A * J), V accumulates the rotations; at convergence columns of A are U * s
    V.assign((size_t)d * d, 0.0);
    for (int i = 0; i < d; i++) V[(size_t)i * d + i] = 1.0;
    for (int sweep = 0; sweep < 60; sweep++) {
        double off = 0;
        for (int p = 0; p < d - 1; p++)
            for (int q = p + 1; q < d; q++) {
                double a = 0, b = 0, c = 0;
                for (int i = 0; i < d; i++) {
                    const double ap = A[(size_t)i * d + p], aq = A[(size_t)i * d + q];
                    a += ap * ap; b += aq * aq; c += ap * aq;
                }
                if (std::fabs(c) <= 1e-15 * std::sqrt(a * b) || c == 0) continue;
                off += c * c / (a * b + 1e-300);
                const double zeta = (b - a) / (2.0 * c);
                const double t = (zeta >= 0 ? 1.0 : -1.0) / (std::fabs(zeta) + std::sqrt(1.0 + zeta * zeta));
                const double cs = 1.0 / std::sqrt(1.0 + t * t), sn = cs * t;
                for (int i = 0; i < d; i++) {
                    const double ap = A[(size_t)i * d + p], aq = A[(size_t)i * d + q];
                    A[(size_t)i * d + p] = cs * ap - sn * aq;
                    A[(size_t)i * d + q] = sn * ap + cs * aq;
                    const double vp = V[(size_t)i * d + p], vq = V[(size_t)i * d + q];
                    V[(size_t)i * d + p] = cs * vp - sn * vq;
                    V[(size_t)i * d + q] = sn * vp + cs * vq;
                }
            }
        if (off < 1e-24) break;
    }
    U.assign((size_t)d * d, 0.0);
    for (int j = 0; j < d; j++) {
        double nrm = 0;
        for (int i = 0; i < d; i++) nrm += A[(size_t)i * d + j] * A[(size_t)i * d + j];
        nrm = std::sqrt(nrm);
        for (int i = 0; i < d; i++) U[(size_t)i * d + j] = nrm > 1e-300 ? A[(size_t)i * d + j] / nrm : (i == j ? 1.0 : 0.0);
    }
}

// OPQMatrix::train (faiss/VectorTransform.cpp:587-744).  r_dev [n][d] residuals (device, not modified).
int opq_train(ivfhnsw_b200_t *h, size_t n, int d, int M, const float *r_dev, int niter, std::vector<float> &A)
{
    cudaStream_t st = h->stream;
    DevBuf xtrain, xproj, recons, codes, colsum, ccov, Adev, cent_dev;
    CU(xtrain.ensure(n * (size_t)d * 4));
    CU(xproj.ensure(n * (size_t)d * 4));
    CU(recons.ensure(n * (size_t)d * 4));
    CU(codes.ensure(n * (size_t)M));
    CU(colsum.ensure((size_t)d * 8));
    CU(ccov.ensure((size_t)d * d * 8));
    CU(Adev.ensure((size_t)d * d * 4));
    CU(cudaMemcpyAsync(xtrain.p, r_dev, n * (size_t)d * 4, cudaMemcpyDeviceToDevice, st));
    CU(cudaMemsetAsync(colsum.p, 0, (size_t)d * 8, st));
    colsum_kernel<<<d, 256, 0, st>>>(n, d, xtrain.as<float>(), colsum.as<double>());
    center_kernel<<<grid_for(n * d), 256, 0, st>>>(n, d, colsum.as<double>(), xtrain.as<float>());
    *LAUNCHES += 2;
    // random rotation: Gram-Schmidt of a Gaussian matrix (float_randn + matrix_qr in the reference)
    A.assign((size_t)d * d, 0.f);
    {
        FaissRng rng(1234);
        std::vector<double> G((size_t)d * d);
        for (size_t i = 0; i < G.size(); i += 2) {
            double a, b, s;
            do {
                a = 2.0 * rng.rand_int() / double(1L << 31) - 1;
                b = 2.0 * rng.rand_int() / double(1L << 31) - 1;
                s = a * a + b * b;
            } while (s >= 1.0 || s == 0.0);
            const double f = std::sqrt(-2.0 * std::log(s) / s);
            G[i] = a * f;
            if (i + 1 < G.size()) G[i + 1] = b * f;
        }
        for (int i = 0; i < d; i++) {
            for (int p = 0; p < i; p++) {
                double dot = 0;
                for (int j = 0; j < d; j++) dot += G[(size_t)i * d + j] * G[(size_t)p * d + j];
                for (int j = 0; j < d; j++) G[(size_t)i * d + j] -= dot * G[(size_t)p * d + j];
            }
            double nrm = 0;
            for (int j = 0; j < d; j++) nrm += G[(size_t)i * d + j] * G[(size_t)i * d + j];
            nrm = std::sqrt(nrm);
            for (int j = 0; j < d; j++) G[(size_t)i * d + j] /= nrm;
        }
        for (size_t i = 0; i < G.size(); i++) A[i] = (float)G[i];
    }
    std::vector<float> cent;
    std::vector<double> C((size_t)d * d), U, V;
    for (int iter = 0; iter < niter; iter++) {
        CU(cudaMemcpyAsync(Adev.p, A.data(), (size_t)d * d * 4, cudaMemcpyHostToDevice, st));
        CU(launch_rotate(st, LAUNCHES, n, d, Adev.as<float>(), false, xtrain.as<float>(), xproj.as<float>()));
        if (int r = pq_kmeans(h, n, d, M, xproj.as<float>(), iter == 0 ? 40 : 4, 1000, iter > 0, cent, cent_dev)) return r;
        CU(launch_pq_encode(st, LAUNCHES, n, d, M, cent_dev.as<float>(), xproj.as<float>(), codes.as<uint8_t>()));
        CU(launch_pq_decode(st, LAUNCHES, n, d, M, cent_dev.as<float>(), codes.as<uint8_t>(), recons.as<float>()));
        dim3 grid((d + 15) / 16, (d + 15) / 16);
        cross_cov_kernel<<<grid, 256, 0, st>>>(n, d, recons.as<float>(), xtrain.as<float>(), ccov.as<double>());
        ++*LAUNCHES;
        CU(cudaMemcpyAsync(C.data(), ccov.p, (size_t)d * d * 8, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        jacobi_svd(d, C, U, V);   // C = recons^T xtrain = U S V^T  ->  A = U V^T
        for (int i = 0; i < d; i++)
            for (int j = 0; j < d; j++) {
                double acc = 0;
                for (int t = 0; t < d; t++) acc += U[(size_t)i * d + t] * V[(size_t)j * d + t];
                A[(size_t)i * d + j] = (float)acc;
            }
    }
    xtrain.release(); xproj.release(); recons.release(); codes.release(); colsum.release(); ccov.release();
    Adev.release(); cent_dev.release();
    return 0;
}

}  // namespace

extern "C" int ivfhnsw_b200_train_pq(ivfhnsw_b200_t *h, size_t n, const float *x)
{
    std::lock_guard<std::mutex> lk(h->mu);
    CU(cudaSetDevice(h->device));
    if (!h->have_centroids) return h->fail(IVFHNSW_B200_ESTATE, "centroids not set (build_quantizer)");
    const int d = (int)h->d, M = (int)h->M;
    cudaStream_t st = h->stream;
    DevBuf xd, lists, sub, alphas, resid, norms;
    CU(xd.ensure(n * (size_t)d * 4));
    CU(lists.ensure(n * 4));
    CU(resid.ensure(n * (size_t)d * 4));
    CU(norms.ensure(n * 4));
    CU(cudaMemcpyAsync(xd.p, x, n * (size_t)d * 4, cudaMemcpyHostToDevice, st));
    // assign train vectors (IndexIVF_HNSW.cpp:539-540)
    if (int r = ivfb200_assign_dev(h, n, xd.as<float>(), lists.as<uint32_t>())) return r;
    const uint8_t *subp = nullptr;
    const float *alp = nullptr;
    if (h->nsubc) {
        // per-group alpha and sub-centroid of every train vector (IndexIVF_HNSW_Grouping.cpp:506-547)
        CU(sub.ensure(n));
        CU(alphas.ensure(h->nc * 4));
        if (int r = ivfb200_group_stage(h, n, xd.as<float>(), lists.as<uint32_t>(), alphas.as<float>(), sub.as<uint8_t>(), 1, nullptr)) return r;
        subp = sub.as<uint8_t>();
        alp = alphas.as<float>();
    }
    // residuals, unrotated
    const bool want_opq = h->do_opq;
    h->do_opq = false;
    int r = ivfb200_encode_dev(h, n, xd.as<float>(), lists.as<uint32_t>(), subp, alp, nullptr, nullptr, nullptr, resid.as<float>());
    h->do_opq = want_opq;
    if (r) return r;
    if (want_opq) {
        // IndexIVF_HNSW.cpp:547-560 (niter 70) / IndexIVF_HNSW_Grouping.cpp:548-563 (niter 100)
        std::vector<float> A;
        if (int rr = opq_train(h, n, d, M, resid.as<float>(), h->nsubc ? 100 : 70, A)) return rr;
        CU(h->opqA.ensure((size_t)d * d * 4));
        CU(cudaMemcpyAsync(h->opqA.p, A.data(), (size_t)d * d * 4, cudaMemcpyHostToDevice, st));
        CU(cudaStreamSynchronize(st));
        h->have_opq = true;
        DevBuf tmp;
        CU(tmp.ensure(n * (size_t)d * 4));
        CU(launch_rotate(st, LAUNCHES, n, d, h->opqA.as<float>(), false, resid.as<float>(), tmp.as<float>()));
        CU(cudaMemcpyAsync(resid.p, tmp.p, n * (size_t)d * 4, cudaMemcpyDeviceToDevice, st));
        CU(cudaStreamSynchronize(st));
        tmp.release();
    }
    // residual PQ (IndexIVF_HNSW.cpp:561-564)
    std::vector<float> cent;
    if (int rr = pq_kmeans(h, n, d, M, resid.as<float>(), 25, 256, false, cent, h->pqc)) return rr;
    h->have_pq = true;
    // norms of the reconstructed train vectors (:566-587), then the 1-D norm PQ (:589-592)
    DevBuf codes;
    CU(codes.ensure(n * (size_t)M));
    if (int rr = ivfb200_encode_dev(h, n, xd.as<float>(), lists.as<uint32_t>(), subp, alp, codes.as<uint8_t>(), nullptr,
                            norms.as<float>(), nullptr))
        return rr;
    std::vector<float> ncent;
    if (int rr = pq_kmeans(h, n, 1, 1, norms.as<float>(), 25, 256, false, ncent, h->norm_table)) return rr;
    h->have_norm = true;
    xd.release(); lists.release(); sub.release(); alphas.release(); resid.release(); norms.release(); codes.release();
    return 0;
}

// copies of the trained model for index->pq / norm_pq / opq_matrix on the host side
extern "C" int ivfhnsw_b200_get_pq(ivfhnsw_b200_t *h, float *pq_centroids, float *norm_centroids, float *opq_A)
{
    CU(cudaSetDevice(h->device));
    cudaStream_t st = h->stream;
    if (pq_centroids) {
        if (!h->have_pq) return h->fail(IVFHNSW_B200_ESTATE, "pq not set");
        CU(cudaMemcpyAsync(pq_centroids, h->pqc.p, h->d * 256 * 4, cudaMemcpyDeviceToHost, st));
    }
    if (norm_centroids) {
        if (!h->have_norm) return h->fail(IVFHNSW_B200_ESTATE, "norm pq not set");
        CU(cudaMemcpyAsync(norm_centroids, h->norm_table.p, 256 * 4, cudaMemcpyDeviceToHost, st));
    }
    if (opq_A) {
        if (!h->have_opq) return h->fail(IVFHNSW_B200_ESTATE, "opq matrix not set");
        CU(cudaMemcpyAsync(opq_A, h->opqA.p, h->d * h->d * 4, cudaMemcpyDeviceToHost, st));
    }
    CU(cudaStreamSynchronize(st));
    return 0;
}

// ------------------------------------------------------------------------------------ coarse k-means
// faiss::Clustering::train (faiss/Clustering.cpp:67-230) for the CENTROID FILE of the index -- the reference leaves it
// to an offline tool ("centroids file create by k-means tool, not here", IndexIVF_HNSW.cpp:703): k = nc clusters of
// the n device vectors x_dev, initial centroids x[perm[i]] (rand_perm with seed + 1, :158-162), niter Lloyd
// iterations whose assignment step is K-A1 itself (the tensor-core GEMM of this library instead of IndexFlatL2) and
// whose update is km_update_centroids (faiss/utils.cpp:1401-1484) including the re-seeding of void clusters with
// faiss' own random stream.  The result becomes the handle's centroid table.
extern "C" int ivfhnsw_b200_train_centroids(ivfhnsw_b200_t *h, size_t n, const float *x_dev, int niter, long seed,
                                            double *objective_out, size_t *nsplit_out)
{
    CU(cudaSetDevice(h->device));
    const size_t k = h->nc;
    const int d = (int)h->d;
    cudaStream_t st = h->stream;
    if (n < k) return h->fail(IVFHNSW_B200_EINVAL, "Number of training points (%zu) should be at least as large as number of clusters (%zu)", n, k);
    if (n >= (1ull << 31)) return h->fail(IVFHNSW_B200_EINVAL, "too many training points");
    if (niter < 0) return h->fail(IVFHNSW_B200_EINVAL, "niter < 0");
    CU(h->centroids.ensure(k * d * 4));
    {   // initial centroids
        std::vector<int> perm;
        faiss_rand_perm(perm, n, seed + 1);
        DevBuf ridx;
        CU(ridx.ensure(k * 4));
        CU(cudaMemcpyAsync(ridx.p, perm.data(), k * 4, cudaMemcpyHostToDevice, st));
        gather_rows_kernel<<<grid_for(k * d), 256, 0, st>>>(k, d, x_dev, ridx.as<int>(), h->centroids.as<float>());
        ++*LAUNCHES;
        CU(cudaStreamSynchronize(st));
    }
    auto centroids_changed = [&]() {
        h->have_centroids = true;
        h->gemm_norms_valid = false;
        h->f16_valid = false;
        if (!h->have_lists) h->have_group_meta = false;
    };
    centroids_changed();
    if (n == k || niter == 0) {
        if (objective_out) *objective_out = 0;
        if (nsplit_out) *nsplit_out = 0;
        return 0;
    }
    DevBuf lab, dist, sums, counts, objb, pairs;
    CU(lab.ensure(n * 4));
    CU(dist.ensure(n * 4));
    CU(sums.ensure(k * d * 8));
    CU(counts.ensure(k * 4));
    CU(objb.ensure(8));
    std::vector<unsigned> hc(k);
    std::vector<uint32_t> hp;
    double obj = 0;
    size_t nsplit_total = 0;
    for (int it = 0; it < niter; it++) {
        // assignment = IndexIVF_HNSW::assign: exact nearest centroid + its distance
        if (int r = ivfb200_coarse_dev(h, n, x_dev, 1, 0, dist.as<float>(), lab.as<uint32_t>())) return r;
        CU(cudaMemsetAsync(sums.p, 0, k * d * 8, st));
        CU(cudaMemsetAsync(counts.p, 0, k * 4, st));
        CU(cudaMemsetAsync(objb.p, 0, 8, st));
        coarse_accum_kernel<<<grid_for(n * d), 256, 0, st>>>(n, d, x_dev, lab.as<uint32_t>(), dist.as<float>(),
                                                           sums.as<double>(), counts.as<unsigned>(), objb.as<double>());
        coarse_mean_kernel<<<grid_for(k * d), 256, 0, st>>>(k, d, sums.as<double>(), counts.as<unsigned>(),
                                                          h->centroids.as<float>());
        *LAUNCHES += 2;
        CU(cudaMemcpyAsync(hc.data(), counts.p, k * 4, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(&obj, objb.p, 8, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        // void clusters: the reference's loop, on the counts (faiss/utils.cpp:1450-1478)
        hp.clear();
        {
            std::vector<size_t> hassign(hc.begin(), hc.end());
            FaissRng rng(1234);
            for (size_t ci = 0; ci < k; ci++) {
                if (hassign[ci] != 0) continue;
                size_t cj;
                for (cj = 0; true; cj = (cj + 1) % k) {
                    const float pr = (hassign[cj] - 1.0f) / (float)(n - k);
                    if (rng.rand_float() < pr) break;
                }
                hp.push_back((uint32_t)ci);
                hp.push_back((uint32_t)cj);
                hassign[ci] = hassign[cj] / 2;
                hassign[cj] -= hassign[ci];
            }
        }
        if (!hp.empty()) {
            CU(pairs.ensure(hp.size() * 4));
            CU(cudaMemcpyAsync(pairs.p, hp.data(), hp.size() * 4, cudaMemcpyHostToDevice, st));
            coarse_split_kernel<<<1, 128, 0, st>>>((int)(hp.size() / 2), d, pairs.as<uint32_t>(), h->centroids.as<float>());
            ++*LAUNCHES;
            CU(cudaStreamSynchronize(st));
            nsplit_total += hp.size() / 2;
        }
        centroids_changed();
    }
    if (objective_out) *objective_out = obj;   // sum of squared distances of the LAST assignment (Clustering.cpp:201-206)
    if (nsplit_out) *nsplit_out = nsplit_total;
    return 0;
}
