// Shared device/host helpers for the ivfhnsw_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cfloat>

#define IVF_FULL_MASK 0xffffffffu
#define IVF_INVALID_SLOT 0xffffffffu

// ADC scan geometry
#define SCAN_THREADS 256
#define SCAN_WARPS 8
#define SCAN_PF 4                          // warp steps (32 codes each) whose loads are in flight together
#define SCAN_SEG 1024                      // list positions per work item
#define SCAN_ITEMS 256                     // work items whose (list, part) is tabulated
#define SCAN_WBUF (SCAN_SEG / 16)          // windows of a work item (a run list is used only when shorter than that)
#define SCAN_WCAP 64                       // per-warp candidate buffer entries

// coarse stage geometry: centroids are processed in groups of COARSE_GROUP, the
// GEMM epilogue keeps the COARSE_T best of each group per query
#define COARSE_GROUP 128
#define COARSE_T 4

// impl 3 of the coarse stage keeps order-preserving integer keys instead of floats:
//   t   = v * S + 1.5 * 2^30      (S = power of two with |v| * S < 2^22 for every (query, centroid) of the batch)
//   key = ((bits(t) - bits(2^30)) << 7) + column          t in [2^30, 2^31) has ulp 128: bits(t) counts quanta
// i.e. key = round(v * S / 128) * 128 + 2^29 + column, < 2^30.  Padding columns get the largest quantum.
#define COARSE_KEY_MAGIC 1610612736.0f             /* 1.5 * 2^30 */
#define COARSE_KEY_BIAS 536870912.0f               /* 2^29: key ~ v * S + 2^29 */
#define COARSE_KEY_INVALID_MIN 0x3fffff80u         /* quantum 2^23 - 1: reserved for padding columns */
static __host__ __device__ __forceinline__ float coarse_key_scale(float qmax2, float cmax2)
{
    // bound of |v| = | ||c||^2 - 2<q,c> | over the batch, then the largest power of two S with bound * S <= 2^22
    const float bound = cmax2 + 2.f * sqrtf(qmax2 * cmax2);
    if (!(bound > 0.f) || !(bound < 1.0e30f)) return 1.f;
    int ex;
    frexpf(bound, &ex);                 // bound = m * 2^ex, m in [0.5, 1)
    return ldexpf(1.f, 22 - ex);
}

static __device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31; }

static __device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel)
{
    uint32_t d;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel));
    return d;
}

// streaming 128-bit / 64-bit loads that do not pollute L1
static __device__ __forceinline__ uint4 ldg_stream_128(const void *p)
{
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
static __device__ __forceinline__ uint2 ldg_stream_64(const void *p)
{
    uint2 r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
    return r;
}
static __device__ __forceinline__ uint32_t ldg_stream_u8(const uint8_t *p)
{
    uint32_t r;
    asm volatile("ld.global.nc.L1::no_allocate.u8 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}

// (distance, sequence) lexicographic "a before b"
static __device__ __forceinline__ bool key_less(float da, uint32_t sa, float db, uint32_t sb)
{
    return da < db || (da == db && sa < sb);
}

// ivfhnsw::fvec_L2sqr (reference utils.cpp:22-51) lane structure: 8 partial sums over
// floor(d/16)*16 dims, partials added left to right.  Executed by 8 consecutive lanes
// (sub = lane & 7); every lane of the octet returns the result.
static __device__ __forceinline__ float l2sqr16_octet(const float *__restrict__ x, const float *__restrict__ y,
                                                      int d, int sub)
{
    const int n = (d >> 4) << 4;
    float acc = 0.f;
    // 16 loads of the (random, usually HBM-resident) row are issued before the first use: the row gather of the
    // coarse re-rank is bound by latency, not by arithmetic (56 % of coarse_select_kernel's stall samples at 2^20 centroids)
    for (int i0 = sub; i0 < n; i0 += 128) {
        float yv[16];
#pragma unroll
        for (int u = 0; u < 16; u++) yv[u] = (i0 + 8 * u < n) ? __ldg(y + i0 + 8 * u) : 0.f;
#pragma unroll
        for (int u = 0; u < 16; u++) {
            if (i0 + 8 * u < n) {
                const float df = x[i0 + 8 * u] - yv[u];
                acc = __fadd_rn(acc, __fmul_rn(df, df));   // unfused, like the pinned C restatement
            }
        }
    }
    const uint32_t lane = lane_id();
    const int base = lane & ~7;
    float r = __shfl_sync(IVF_FULL_MASK, acc, base);
#pragma unroll
    for (int l = 1; l < 8; l++) r += __shfl_sync(IVF_FULL_MASK, acc, base + l);
    return r;
}
