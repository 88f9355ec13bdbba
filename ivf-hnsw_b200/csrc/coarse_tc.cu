// Coarse GEMM on the 5th-gen tensor cores (tcgen05 + TMEM + TMA), impl 1 of launch_coarse.
// (kernel body lands in a later commit of this round; the conversion helpers are already used)
#include <cuda_bf16.h>

#include "common.cuh"
#include "kernels.h"

namespace ivfb200 {

__global__ void f32_to_bf16_kernel(size_t n4, const float4 *__restrict__ x, uint2 *__restrict__ out)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        const float4 v = x[i];
        const __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
        uint2 o;
        o.x = *reinterpret_cast<const uint32_t *>(&a);
        o.y = *reinterpret_cast<const uint32_t *>(&b);
        out[i] = o;
    }
}

cudaError_t launch_f32_to_bf16(cudaStream_t st, uint64_t *launches, size_t n, const float *x, void *out)
{
    if (n == 0) return cudaSuccess;
    const size_t n4 = n / 4;
    size_t blocks = (n4 + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    f32_to_bf16_kernel<<<(unsigned)blocks, 256, 0, st>>>(n4, reinterpret_cast<const float4 *>(x),
                                                       reinterpret_cast<uint2 *>(out));
    ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_coarse_tc(cudaStream_t st, uint64_t *launches, int nq, int nc, int d, const float *queries,
                             const void *cent_bf16, void *xq_bf16, const float *cnorm_gemm, int self_off,
                             float *grp_val, uint32_t *grp_idx)
{
    (void)st; (void)launches; (void)nq; (void)nc; (void)d; (void)queries; (void)cent_bf16; (void)xq_bf16;
    (void)cnorm_gemm; (void)self_off; (void)grp_val; (void)grp_idx;
    return cudaErrorNotSupported;
}

}  // namespace ivfb200
