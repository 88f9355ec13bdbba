// ivfhnsw_b200_coarse_noalloc_dev: the exact coarse search (K-A2, fp32 SIMT GEMM + select + re-rank of
// coarse.cu) for callers that must not allocate: CUDA-graph capture, tight serving loops, harnesses that forbid
// cudaMalloc.  It has NO counterpart in the reference.
//
// The regular entry point (ivfhnsw_b200_coarse_dev, api.cu) keeps the GEMM's per-(query, group) kept values in a
// cudaMalloc'ed workspace of the handle.  Here the same intermediates live in ONE file-scope device array of
// IVFHNSW_B200_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit.  Consequences:
//   * sizes whose scratch does not fit are rejected with EINVAL (no fallback to an allocating path);
//   * every call on a device shares the array and finds the previous call's contents in it: the one word that must
//     start at zero (the select stage's sticky overflow flags) is cleared on the stream first, everything else is
//     fully overwritten before it is read;
//   * two calls that overlap on the GPU (different streams / handles on one device) RACE on the array: callers
//     serialise them.
// The array's name is only a host-side stand-in in host code; its device address is taken once per device with
// cudaGetSymbolAddress (with that device current) and that pointer is what the kernels get.
#include <cuda_runtime.h>

#include <mutex>

#include "common.cuh"
#include "handle.h"

namespace {

static __device__ __align__(256) unsigned char g_noalloc_scratch[IVFHNSW_B200_NOALLOC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
std::mutex g_scratch_mu;
void *g_scratch_ptr[kMaxDevices] = {nullptr};

// device address of the scratch array on `device`, which must be the CURRENT device
cudaError_t scratch_address(int device, void **out)
{
    if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lk(g_scratch_mu);
    if (!g_scratch_ptr[device]) {
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        g_scratch_ptr[device] = p;
    }
    *out = g_scratch_ptr[device];
    return cudaSuccess;
}

size_t up256(size_t x) { return (x + 255) & ~(size_t)255; }

}  // namespace

using namespace ivfb200;

extern "C" int ivfhnsw_b200_coarse_noalloc_dev(ivfhnsw_b200_t *h, size_t nq, const float *xq_dev, size_t nprobe,
                                               float *dists_dev, uint32_t *idxs_dev, uint32_t *flags_dev)
{
    if (!h) return IVFHNSW_B200_EINVAL;
    CU(cudaSetDevice(h->device));
    if (!h->have_centroids) return h->fail(IVFHNSW_B200_ESTATE, "centroids not set (build_quantizer)");
    if (nprobe == 0 || nprobe > h->nc || nprobe > 2048) return h->fail(IVFHNSW_B200_EINVAL, "nprobe %zu out of range", nprobe);
    if (nq == 0) return 0;
    if (!xq_dev || !dists_dev || !idxs_dev) return h->fail(IVFHNSW_B200_EINVAL, "null device pointer");
    const int slack = (int)h->coarse_slack;
    int cand_cap = 0;
    (void)coarse_work_bytes(1, (int)h->nc, (int)nprobe + slack, &cand_cap);
    const size_t ngroups = (h->nc + COARSE_GROUP - 1) / COARSE_GROUP;
    // scratch layout: [flags 256 B][||c||^2 nc floats][kept values nq*ngroups*T floats][kept ids, same]
    const size_t off_norm = 256;
    const size_t off_val = off_norm + up256(h->nc * 4);
    const size_t off_idx = off_val + up256(nq * ngroups * COARSE_T * 4);
    const size_t need = off_idx + up256(nq * ngroups * COARSE_T * 4);
    if (need > (size_t)IVFHNSW_B200_NOALLOC_SCRATCH_BYTES)
        return h->fail(IVFHNSW_B200_EINVAL,
                       "coarse_noalloc: nq=%zu x nc=%zu needs %zu B of scratch, the static array holds %zu B "
                       "(nq * ceil(nc/128) * 32 + nc * 4 + 768 <= %zu): split the batch or use ivfhnsw_b200_coarse_dev",
                       nq, h->nc, need, (size_t)IVFHNSW_B200_NOALLOC_SCRATCH_BYTES,
                       (size_t)IVFHNSW_B200_NOALLOC_SCRATCH_BYTES);
    void *base = nullptr;
    CU(scratch_address(h->device, &base));
    unsigned char *s = static_cast<unsigned char *>(base);
    cudaStream_t st = h->stream;
    uint32_t *flags = reinterpret_cast<uint32_t *>(s);
    float *cnorm = reinterpret_cast<float *>(s + off_norm);
    CU(cudaMemsetAsync(flags, 0, 256, st));   // the only part of the shared array that must start at zero
    CU(launch_row_norms(st, LAUNCHES, h->nc, (int)h->d, h->centroids.as<float>(), cnorm, coarse_dims((int)h->d)));
    CoarseWork W;
    W.grp_val = reinterpret_cast<float *>(s + off_val);
    W.grp_idx = reinterpret_cast<uint32_t *>(s + off_idx);
    W.cand = nullptr;
    W.cand_d = nullptr;
    W.err = flags;
    W.cand_cap = cand_cap;
    TcSide S{};   // the tensor-core operands are not used by impl 0 (all pointers null)
    CU(launch_coarse(st, LAUNCHES, /*impl=*/0, (int)nq, (int)h->nc, (int)h->d, (int)nprobe, slack, xq_dev,
                     h->centroids.as<float>(), cnorm, S, W, -1, dists_dev, idxs_dev, nullptr));
    if (flags_dev) CU(cudaMemcpyAsync(flags_dev, flags, 4, cudaMemcpyDeviceToDevice, st));
    h->stats.coarse_flops += (uint64_t)2 * h->nc * h->d * nq;
    return 0;
}
