// micro-benchmark: random row gather from an L2-resident table, thread-per-row sector loads vs warp-coalesced rows
#include <cstdio>
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>
template <int ROWB>
__global__ void __launch_bounds__(512) g_thread_row(const uint8_t *table, const uint32_t *idx, int per_thread, float *out)
{
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    float acc = 0.f;
    for (int g = 0; g < per_thread; g++) {
        const float *r = reinterpret_cast<const float *>(table + (size_t)idx[(size_t)g * gridDim.x * blockDim.x + t] * ROWB);
#pragma unroll 4
        for (int i = 0; i < ROWB / 32; i++) {
            float v[8];
            asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]) : "l"(r + i * 8));
#pragma unroll
            for (int j = 0; j < 8; j++) acc += v[j];
        }
    }
    if (acc == 12345.f) out[0] = acc;
}
// a warp reads whole rows: lane l takes bytes [16 l, 16 l + 16) (+512 for longer rows); UNR rows in flight
template <int ROWB, int UNR>
__global__ void __launch_bounds__(512) g_warp_row(const uint8_t *table, const uint32_t *idx, int per_warp, float *out)
{
    const int lane = threadIdx.x & 31;
    const size_t w = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t *my = idx + w * per_warp;
    float acc = 0.f;
    for (int g = 0; g < per_warp; g += 32) {
        const uint32_t mine = my[g + lane];
#pragma unroll
        for (int u0 = 0; u0 < 32; u0 += UNR) {
            float4 v[UNR];
#pragma unroll
            for (int u = 0; u < UNR; u++) {
                const uint32_t row = __shfl_sync(0xffffffffu, mine, u0 + u);
                v[u] = make_float4(0, 0, 0, 0);
                if (lane * 16 < ROWB) v[u] = __ldg(reinterpret_cast<const float4 *>(table + (size_t)row * ROWB) + lane);
            }
#pragma unroll
            for (int u = 0; u < UNR; u++) acc += v[u].x + v[u].y + v[u].z + v[u].w;
        }
    }
    if (acc == 12345.f) out[0] = acc;
}
template <int ROWB>
void run(int nrows)
{
    const int blocks = 148 * 4, threads = 512, per_thread = 16;
    const size_t nidx = (size_t)blocks * threads * per_thread;
    uint8_t *table; uint32_t *idx; float *out;
    cudaMalloc(&table, (size_t)nrows * ROWB); cudaMemset(table, 0, (size_t)nrows * ROWB);
    std::vector<uint32_t> h(nidx);
    uint32_t x = 12345;
    for (auto &v : h) { x = x * 1664525u + 1013904223u; v = (x >> 8) % nrows; }
    cudaMalloc(&idx, nidx * 4); cudaMemcpy(idx, h.data(), nidx * 4, cudaMemcpyHostToDevice);
    cudaMalloc(&out, 4);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    float ms;
    for (int it = 0; it < 2; it++) {
        cudaEventRecord(a);
        g_thread_row<ROWB><<<blocks, threads>>>(table, idx, per_thread, out);
        cudaEventRecord(b); cudaEventSynchronize(b); cudaEventElapsedTime(&ms, a, b);
        printf("row %d B, table %d MB: thread-per-row LDG.256 : %.3f ms %.2f Grows/s %.0f GB/s (%s)\n", ROWB, (int)((size_t)nrows * ROWB >> 20), ms, nidx / ms / 1e6, nidx * (double)ROWB / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
        cudaEventRecord(a);
        g_warp_row<ROWB, 4><<<blocks, threads>>>(table, idx, per_thread * 32, out);
        cudaEventRecord(b); cudaEventSynchronize(b); cudaEventElapsedTime(&ms, a, b);
        printf("row %d B, table %d MB: warp-per-row LDG.128 x4  : %.3f ms %.2f Grows/s %.0f GB/s (%s)\n", ROWB, (int)((size_t)nrows * ROWB >> 20), ms, nidx / ms / 1e6, nidx * (double)ROWB / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
        cudaEventRecord(a);
        g_warp_row<ROWB, 8><<<blocks, threads>>>(table, idx, per_thread * 32, out);
        cudaEventRecord(b); cudaEventSynchronize(b); cudaEventElapsedTime(&ms, a, b);
        printf("row %d B, table %d MB: warp-per-row LDG.128 x8  : %.3f ms %.2f Grows/s %.0f GB/s (%s)\n", ROWB, (int)((size_t)nrows * ROWB >> 20), ms, nidx / ms / 1e6, nidx * (double)ROWB / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
    }
    cudaFree(table); cudaFree(idx); cudaFree(out);
}
int main()
{
    run<384>(65536);      // C2: 25 MB table
    run<512>(131072);     // c4s: 64 MB table
    run<512>(1048576);    // C4: 512 MB table (HBM)
    return 0;
}
