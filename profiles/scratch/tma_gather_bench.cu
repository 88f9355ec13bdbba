// micro-benchmark: how fast can cp.async.bulk gather 384-byte rows from an L2-resident table into smem?
#include <cstdio>
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>
#define ROWB 384
#define STAGES 2
#define WARPS 8
__device__ __forceinline__ uint32_t s32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void __launch_bounds__(WARPS * 32) gather(const uint8_t *table, const uint32_t *idx, int per_warp, float *out)
{
    extern __shared__ __align__(128) uint8_t sm[];
    __shared__ uint64_t bars[WARPS * STAGES];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint8_t *buf = sm + (size_t)warp * STAGES * 32 * ROWB;
    if (lane == 0)
        for (int s = 0; s < STAGES; s++) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bars[warp * STAGES + s])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncwarp();
    const uint32_t *my = idx + ((size_t)blockIdx.x * WARPS + warp) * per_warp;
    float acc = 0.f;
    auto issue = [&](int g, int s) {
        const uint32_t bar = s32(&bars[warp * STAGES + s]);
        if (lane == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(32 * ROWB) : "memory");
        __syncwarp();
        const uint32_t row = my[g * 32 + lane];
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(s32(buf + ((size_t)s * 32 + lane) * ROWB)), "l"(table + (size_t)row * ROWB), "r"(ROWB), "r"(bar) : "memory");
    };
    const int ngroups = per_warp / 32;
    issue(0, 0);
    for (int g = 0; g < ngroups; g++) {
        const int s = g & 1;
        if (g + 1 < ngroups) issue(g + 1, s ^ 1);
        const uint32_t bar = s32(&bars[warp * STAGES + s]);
        uint32_t done = 0, par = (g >> 1) & 1;
        while (!done)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0,1,0,p; }" : "=r"(done) : "r"(bar), "r"(par) : "memory");
        const float4 *r = reinterpret_cast<const float4 *>(buf + ((size_t)s * 32 + lane) * ROWB);
#pragma unroll
        for (int i = 0; i < ROWB / 16; i++) { const float4 v = r[i]; acc += v.x + v.y + v.z + v.w; }
        __syncwarp();
    }
    if (acc == 12345.f) out[0] = acc;
}
__global__ void __launch_bounds__(WARPS * 32) gather_ldg(const uint8_t *table, const uint32_t *idx, int per_warp, float *out)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t *my = idx + ((size_t)blockIdx.x * WARPS + warp) * per_warp;
    float acc = 0.f;
    for (int g = 0; g < per_warp / 32; g++) {
        const float4 *r = reinterpret_cast<const float4 *>(table + (size_t)my[g * 32 + lane] * ROWB);
#pragma unroll 8
        for (int i = 0; i < ROWB / 16; i++) { const float4 v = __ldg(r + i); acc += v.x + v.y + v.z + v.w; }
    }
    if (acc == 12345.f) out[0] = acc;
}
int main()
{
    const int nrows = 65536, per_warp = 32 * 64, blocks = 148 * 6;
    uint8_t *table; uint32_t *idx; float *out;
    cudaMalloc(&table, (size_t)nrows * ROWB); cudaMemset(table, 0, (size_t)nrows * ROWB);
    const size_t nidx = (size_t)blocks * WARPS * per_warp;
    std::vector<uint32_t> h(nidx);
    uint32_t x = 12345;
    for (auto &v : h) { x = x * 1664525u + 1013904223u; v = (x >> 8) % nrows; }
    cudaMalloc(&idx, nidx * 4); cudaMemcpy(idx, h.data(), nidx * 4, cudaMemcpyHostToDevice);
    cudaMalloc(&out, 4);
    const size_t smem = (size_t)WARPS * STAGES * 32 * ROWB;
    cudaFuncSetAttribute(gather, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int it = 0; it < 2; it++) {
        cudaEventRecord(a);
        gather<<<blocks, WARPS * 32, smem>>>(table, idx, per_warp, out);
        cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        printf("TMA bulk gather: %.3f ms, %.1f Mrows/s, %.1f GB/s (%s)\n", ms, nidx / ms / 1e3, nidx * (double)ROWB / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
        cudaEventRecord(a);
        gather_ldg<<<blocks, WARPS * 32>>>(table, idx, per_warp, out);
        cudaEventRecord(b); cudaEventSynchronize(b);
        cudaEventElapsedTime(&ms, a, b);
        printf("LDG.128 thread-per-row: %.3f ms, %.1f Mrows/s, %.1f GB/s (%s)\n", ms, nidx / ms / 1e3, nidx * (double)ROWB / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
