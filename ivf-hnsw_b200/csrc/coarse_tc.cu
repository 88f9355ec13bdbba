// Coarse GEMM on the 5th-gen tensor cores (tcgen05 + TMEM + TMA), impl 1 of launch_coarse.
//
// value(q, c) = ||c||^2 - 2 <q, c> for every (query, centroid), reduced on the fly to the
// COARSE_T smallest values of every group of COARSE_GROUP centroids (the same contract as the
// fp32 SIMT kernel in coarse.cu, so the select / verify / exact re-rank stage is shared).
//
// Exactness on fp16 tensor cores: every fp32 operand row is split into two fp16 rows
//     x * s = hi + lo        (s a power of two, |x * s| < 2^14, lo = the fp16-rounded remainder)
// and the kernel accumulates  hi.hi + lo.hi + hi.lo  in the fp32 TMEM accumulator.  The dropped
// lo.lo term and the rounding of lo are ~2^-22 relative, i.e. fp32-class accuracy, at three
// MMAs per (k-step).  nterms == 1 (the default, coarse_impl 2) leaves out hi.lo, i.e. q.c_lo: only the centroids'
// hi blocks are streamed and the select stage widens its threshold by twice the bound 2^-10 ||q|| max||c|| of what
// that drops (coarse.cu), so the answer is still the exact top-nprobe.
//
// CTA = 12 warps, persistent over work units (a tile of 128*MT queries x a range of centroid groups):
//   warp 0      TMA producer: the query operand (hi and lo, resident per unit) and a ring of
//               16 KB centroid blocks (128 centroids x 64 dims, hi or lo), 128B-swizzled
//   warp 1      MMA issuer: tcgen05.mma cta_group::1 kind::f16, M=128 N=128 K=16, accumulators
//               in TMEM, double buffered (2 stages x MT x 128 columns = all 512 columns for MT=2)
//   warp 2      TMEM allocator
//   warps 4-11  epilogue: thread <-> query row, tcgen05.ld 32 columns at a time, scale, add
//               ||c||^2, pack the column index into the low 7 mantissa bits and run a
//               branch-free sorted insert (7 FMNMX) into the running 4-best of the group
#include <cuda.h>
#include <cuda_fp16.h>

#include <cstdlib>

#include "common.cuh"
#include "kernels.h"

namespace ivfb200 {

// --------------------------------------------------------------------------------------------- operand split
__global__ void maxabs_kernel(size_t n, const float *__restrict__ x, uint32_t *__restrict__ out_bits)
{
    float m = 0.f;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float v = fabsf(x[i]);
        if (v < __int_as_float(0x7f800000)) m = fmaxf(m, v);
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) m = fmaxf(m, __shfl_xor_sync(IVF_FULL_MASK, m, off));
    if ((threadIdx.x & 31) == 0) atomicMax(out_bits, __float_as_uint(m));
}

static __device__ __forceinline__ float pow2_scale_for(float maxabs)
{
    if (!(maxabs > 0.f)) return 1.f;
    int ex;
    frexpf(maxabs, &ex);                // maxabs = m * 2^ex, m in [0.5, 1)
    return ldexpf(1.f, 14 - ex);        // maxabs * scale in [2^13, 2^14)
}

// one warp per row.  shared_max != nullptr: all rows use that global scale (centroids);
// otherwise each row has its own scale and fac[row] = -2 / (s_row * s_other) is written.
__global__ void split_rows_kernel(size_t n, int d, int dp, const float *__restrict__ x,
                                  const uint32_t *__restrict__ shared_max, const uint32_t *__restrict__ other_max,
                                  __half *__restrict__ hi, __half *__restrict__ lo, float *__restrict__ fac,
                                  uint32_t *__restrict__ stat /* [2]: max ||row||^2 bits, any non-zero lo; may be null */,
                                  int dk /* dims the GEMM uses */, uint32_t idmul = 0 /* output row j = input row j*idmul mod n */)
{
    const int lane = threadIdx.x & 31;
    const size_t row = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= n) return;
    const float *xr = x + (idmul ? (size_t)(((uint64_t)row * idmul) % n) : row) * d;
    float s;
    if (shared_max) s = pow2_scale_for(__uint_as_float(*shared_max));
    else {
        float m = 0.f;
        for (int i = lane; i < d; i += 32) {
            const float v = fabsf(xr[i]);
            if (v < __int_as_float(0x7f800000)) m = fmaxf(m, v);
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) m = fmaxf(m, __shfl_xor_sync(IVF_FULL_MASK, m, off));
        s = pow2_scale_for(m);
    }
    float n2 = 0.f;
    bool any_lo = false;
    for (int i = lane; i < dp; i += 32) {
        float v = i < d ? xr[i] * s : 0.f;
        const __half h = __float2half_rn(v);
        const __half l = __float2half_rn(v - __half2float(h));
        hi[row * dp + i] = h;
        lo[row * dp + i] = l;
        if (i < dk) {
            n2 = fmaf(xr[i], xr[i], n2);
            any_lo |= (v - __half2float(h)) != 0.f;
        }
    }
    if (stat) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) n2 += __shfl_xor_sync(IVF_FULL_MASK, n2, off);
        const bool any = __any_sync(IVF_FULL_MASK, any_lo);
        if (lane == 0) {
            if (n2 < 1.0e30f) atomicMax(stat, __float_as_uint(n2 * 1.0001f));   // sums differ by rounding: head room
            if (any) atomicOr(stat + 1, 1u);
        }
    }
    if (fac && lane == 0) {
        const float so = pow2_scale_for(__uint_as_float(*other_max));
        fac[row] = -2.f / (s * so);
    }
}

__global__ void permute_norms_kernel(size_t n, uint32_t idmul, const float *__restrict__ in, float *__restrict__ out)
{
    const size_t j = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j < n) out[j] = in[idmul ? (size_t)(((uint64_t)j * idmul) % n) : j];
}

// id(col) = col * A mod nc with A the inverse of G ~ nc / golden ratio: consecutive ids sit G columns apart (mod nc), the
// most evenly spread arrangement there is (three-distance theorem), so a run of neighbouring ids never piles up in one
// group of 128 columns
uint32_t coarse_col_multiplier(size_t nc)
{
    if (nc < 4) return 0;
    auto gcd = [](uint64_t a, uint64_t b) { while (b) { const uint64_t t = a % b; a = b; b = t; } return a; };
    uint64_t G = (uint64_t)((double)nc * 0.6180339887498949);
    if (G < 1) G = 1;
    while (gcd(G, nc) != 1) G++;
    // modular inverse of G (extended Euclid)
    int64_t t0 = 0, t1 = 1, r0 = (int64_t)nc, r1 = (int64_t)(G % nc);
    while (r1) {
        const int64_t qv = r0 / r1;
        const int64_t t2 = t0 - qv * t1; t0 = t1; t1 = t2;
        const int64_t r2 = r0 - qv * r1; r0 = r1; r1 = r2;
    }
    if (t0 < 0) t0 += (int64_t)nc;
    return (uint32_t)t0;
}

cudaError_t launch_tc_prepare_centroids(cudaStream_t st, uint64_t *launches, size_t nc, int d, const float *centroids,
                                        const float *cnorm_gemm, const TcSide &S)
{
    cudaError_t e = cudaMemsetAsync(S.cmax, 0, 8, st);
    if (e != cudaSuccess) return e;
    size_t blocks = (nc * d + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    maxabs_kernel<<<(unsigned)blocks, 256, 0, st>>>(nc * d, centroids, S.cmax);
    ++*launches;
    // cmax[1] = bits of max ||c||^2: the select margin of the two-product mode
    maxabs_kernel<<<(unsigned)((nc + 255) / 256 < 1184 ? (nc + 255) / 256 : 1184), 256, 0, st>>>(nc, cnorm_gemm, S.cmax + 1);
    ++*launches;
    split_rows_kernel<<<(unsigned)((nc + 7) / 8), 256, 0, st>>>(nc, d, S.dp, centroids, S.cmax, nullptr, S.chi, S.clo,
                                                              nullptr, nullptr, 0, S.idmul);
    ++*launches;
    permute_norms_kernel<<<(unsigned)((nc + 255) / 256), 256, 0, st>>>(nc, S.idmul, cnorm_gemm, S.cnp);
    ++*launches;
    return cudaGetLastError();
}

// --------------------------------------------------------------------------------------------- PTX helpers
static __device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

static __device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
static __device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
static __device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// bounded wait: a protocol bug traps (-> launch error) instead of hanging the device.  try_wait suspends
// the thread in hardware (up to the hint) instead of burning issue slots the epilogue warps need.
static __device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    uint32_t done = 0;
    for (uint32_t spins = 0;; spins++) {
        asm volatile(
            "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\tselp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(bar), "r"(parity), "r"(20000u)
            : "memory");
        if (done) break;
        if (spins > (1u << 22)) __trap();
    }
}
// one lane of a converged warp (the others get false)
static __device__ __forceinline__ bool elect_one_sync()
{
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
static __device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *map, int x, int y, uint32_t bar)
{
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
        ::"r"(dst), "l"(map), "r"(x), "r"(y), "r"(bar)
        : "memory");
}
static __device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                                uint32_t accumulate)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
static __device__ __forceinline__ void umma_commit(uint32_t bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
static __device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
static __device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
static __device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

static __device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32])
{
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}

// K-major, 128B-swizzled operand tile: rows of 128 B, 8-row groups 1024 B apart
static __device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr)
{
    return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}

#define TC_THREADS 384
#define TC_BLOCK_BYTES 16384          // 128 rows x 64 fp16
#define TC_INVALID 3.0e38f            // masked column; stays finite with index bits packed in

struct TcParams {
    int nq, nc, d, db;                // db = ceil(d / 64) column blocks
    int ntile_c, csplit, tiles_per_unit, nunits, nstages;
    int nterms;                       // centroid operand terms: 2 = hi and lo (three products), 1 = hi only (two products)
    int qterms;                       // query operand terms: 2 = hi and lo, 1 = hi only (impl 3)
    const uint32_t *cmax;             // [2] bits of max |c component|, max ||c||^2   (integer keys: scale)
    const uint32_t *qstat;            // [2] bits of max ||q||^2 of the batch, any-lo flag
    int gshift;                       // integer keys: 2 = the kept keys of 4 consecutive tiles are merged into ONE record of
                                      // 4 (a 512-centroid super-group): a quarter of the output volume for the select stage
    const float *cnorm;               // [nc] ||c||^2
    const float *qfac;                // [nq] -2 / (s_q * s_c)
    float *grp_val;                   // [nq][ntile_c][COARSE_T] packed (value | column index)
    uint32_t keep_mask;               // ~127, passed at run time so that packing is ONE lop3 (reg & reg | imm)
    uint32_t key_mul;                 // 128, passed at run time so that the key is built by ONE IMAD (FMA pipe): a
                                      // literal power of two becomes a LEA on the ALU pipe, which is the busy one
};

// KEEP: values kept per (query, group of 128 centroids): 4, or 3 / 2 when the groups are many (see launch_coarse)
// IKEY: integer keys (impl 3, common.cuh): one FFMA + one IMAD on the FMA pipe build the key, KEEP * 2 - 1 integer
// min/max insert it -- the ALU pipe (half rate for min/max and logic ops) is what bounds this kernel
// PRE (IKEY only): 2 = the keys of columns (2j, 2j+1) are reduced to their minimum BEFORE the sorted insert (1 + (2 KEEP
// - 1) min/max per two elements instead of 2 (2 KEEP - 1)); a kept key then stands for its column PAIR and the select
// stage scores the other column of every surviving pair exactly
template <int MT, int KEEP, bool IKEY, int DIAG = 0, int PRE = 1>
__global__ void __launch_bounds__(TC_THREADS, 1)
coarse_tc_kernel(const __grid_constant__ CUtensorMap tm_qhi, const __grid_constant__ CUtensorMap tm_qlo,
                 const __grid_constant__ CUtensorMap tm_chi, const __grid_constant__ CUtensorMap tm_clo, TcParams p)
{
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // 128B swizzle wants 1024 B
    const int db = p.db, nst = p.nstages;
    uint8_t *sA = smem;                                            // [MT][db][2] blocks
    uint8_t *sB = sA + (size_t)MT * db * 2 * TC_BLOCK_BYTES;       // [nst] blocks
    float *s_cn = reinterpret_cast<float *>(sB + (size_t)nst * TC_BLOCK_BYTES);   // [MT][2][128]
    uint64_t *bars = reinterpret_cast<uint64_t *>(s_cn + MT * 2 * 128);
    uint32_t *s_tmem = reinterpret_cast<uint32_t *>(bars + 32);
    // barrier map: [0,nst) full, [nst,2nst) empty, then a_full, a_empty, tmem_full[2], tmem_empty[2]
    const uint32_t bar0 = smem_u32(bars);
    auto FULL = [&](int s) { return bar0 + 8u * s; };
    auto EMPTY = [&](int s) { return bar0 + 8u * (nst + s); };
    const uint32_t A_FULL = bar0 + 8u * (2 * nst), A_EMPTY = A_FULL + 8;
    auto T_FULL = [&](int s) { return A_FULL + 16u + 8u * s; };
    auto T_EMPTY = [&](int s) { return A_FULL + 32u + 8u * s; };

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    constexpr int EPI_WARPS = 4 * MT;

    if (warp == 1 && lane == 0) {
        for (int s = 0; s < nst; s++) { mbar_init(FULL(s), 1); mbar_init(EMPTY(s), 1); }
        mbar_init(A_FULL, 1);
        mbar_init(A_EMPTY, 1);
        for (int s = 0; s < 2; s++) { mbar_init(T_FULL(s), 1); mbar_init(T_EMPTY(s), EPI_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(s_tmem)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *s_tmem;

    const int d = p.d;
    if (warp == 0) {
        // ================================================================ TMA producer
        // (one lane runs the whole loop: two loads per tile, nowhere near the issue limit -- the converged-warp form
        // of the MMA issuer below was measured slower here: 32 polling lanes take issue slots from the epilogue)
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0, a_phase = 0;
            for (int u = blockIdx.x; u < p.nunits; u += gridDim.x) {
                const int qt = u / p.csplit, cs = u - qt * p.csplit;
                const int t0 = cs * p.tiles_per_unit;
                const int t1 = min(t0 + p.tiles_per_unit, p.ntile_c);
                const int q0 = qt * 128 * MT;
                mbar_wait(A_EMPTY, a_phase ^ 1);
                mbar_expect_tx(A_FULL, (uint32_t)(MT * db * p.qterms * TC_BLOCK_BYTES));
                for (int mt = 0; mt < MT; mt++)
                    for (int j = 0; j < db; j++) {
                        const uint32_t dst = smem_u32(sA + (size_t)((mt * db + j) * 2) * TC_BLOCK_BYTES);
                        tma_load_2d(dst, &tm_qhi, j * 64, q0 + mt * 128, A_FULL);
                        if (p.qterms == 2) tma_load_2d(dst + TC_BLOCK_BYTES, &tm_qlo, j * 64, q0 + mt * 128, A_FULL);
                    }
                a_phase ^= 1;
                for (int t = t0; t < t1; t++) {
                    for (int j = 0; j < db; j++) {
#pragma unroll
                        for (int h = 0; h < p.nterms; h++) {
                            mbar_wait(EMPTY(stage), phase ^ 1);
                            mbar_expect_tx(FULL(stage), TC_BLOCK_BYTES);
                            tma_load_2d(smem_u32(sB + (size_t)stage * TC_BLOCK_BYTES), h ? &tm_clo : &tm_chi, j * 64,
                                        t * 128, FULL(stage));
                            if (++stage == nst) { stage = 0; phase ^= 1; }
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ================================================================ MMA issuer
        // The WHOLE warp runs the control flow (barrier waits, loop counters) and one elected lane issues the
        // tcgen05.mma instructions.  With the loop inside an `if (lane == 0)` the compiler wraps every tcgen05.mma in
        // an election loop and rebuilds both 64-bit descriptors from scratch: ~45 dependent instructions = ~200 clocks
        // per 64-clock MMA, which made the ISSUE rate, not the tensor pipe, the ceiling of this kernel (measured:
        // 369 ms per 1M x 2^20 batch with the epilogue's arithmetic removed, 3x the tensor pipe's time).
        // kind::f16: D fp32, A/B fp16 K-major, M = 128, N = 128
        const uint32_t idesc = (1u << 4) | ((128u >> 3) << 17) | ((128u >> 4) << 24);
        const bool leader = elect_one_sync();
        // descriptors of the resident query operand: one per (M tile, column block, hi/lo); a k-step of 16 fp16 = 32 B
        // advances the 16-byte-granular start address field by 2
        const uint64_t adesc0 = umma_desc_sw128(smem_u32(sA));
        const uint64_t bdesc0 = umma_desc_sw128(smem_u32(sB));
        constexpr uint64_t BLK = TC_BLOCK_BYTES >> 4;      // one 16 KB block in descriptor address units
        int stage = 0, as = 0;
        uint32_t phase = 0, a_phase = 0, t_phase = 0;
        for (int u = blockIdx.x; u < p.nunits; u += gridDim.x) {
            const int cs = u % p.csplit;
            const int t0 = cs * p.tiles_per_unit;
            const int t1 = min(t0 + p.tiles_per_unit, p.ntile_c);
            mbar_wait(A_FULL, a_phase);
            a_phase ^= 1;
            for (int t = t0; t < t1; t++) {
                mbar_wait(T_EMPTY(as), t_phase ^ 1);
                tc_fence_after();
                for (int j = 0; j < db; j++) {
                    const int ks = min(4, (d - j * 64 + 15) >> 4);
                    // centroid hi block: q_hi.c_hi (and q_lo.c_hi)
                    mbar_wait(FULL(stage), phase);
                    tc_fence_after();
                    if (leader) {
                        const uint64_t bd = bdesc0 + (uint64_t)stage * BLK;
#pragma unroll
                        for (int mt = 0; mt < MT; mt++) {
                            const uint64_t ad = adesc0 + (uint64_t)((mt * db + j) * 2) * BLK;
                            const uint32_t dcol = tmem_base + (uint32_t)(as * MT * 128 + mt * 128);
#pragma unroll
                            for (int k = 0; k < 4; k++) {
                                if ((DIAG & 1) || k >= ks) continue;      // DIAG: timing experiment, no tensor-core work
                                umma_f16(dcol, ad + 2u * k, bd + 2u * k, idesc, (j | k) ? 1u : 0u);
                                if (p.qterms == 2) umma_f16(dcol, ad + BLK + 2u * k, bd + 2u * k, idesc, 1u);
                            }
                        }
                        umma_commit(EMPTY(stage));
                    }
                    __syncwarp();
                    if (++stage == nst) { stage = 0; phase ^= 1; }
                    if (p.nterms < 2) continue;   // the q.c_lo product is covered by the select margin
                    // centroid lo block: q_hi.c_lo
                    mbar_wait(FULL(stage), phase);
                    tc_fence_after();
                    if (leader) {
                        const uint64_t bd = bdesc0 + (uint64_t)stage * BLK;
#pragma unroll
                        for (int mt = 0; mt < MT; mt++) {
                            const uint64_t ad = adesc0 + (uint64_t)((mt * db + j) * 2) * BLK;
                            const uint32_t dcol = tmem_base + (uint32_t)(as * MT * 128 + mt * 128);
#pragma unroll
                            for (int k = 0; k < 4; k++)
                                if (k < ks) umma_f16(dcol, ad + 2u * k, bd + 2u * k, idesc, 1u);
                        }
                        umma_commit(EMPTY(stage));
                    }
                    __syncwarp();
                    if (++stage == nst) { stage = 0; phase ^= 1; }
                }
                if (leader) umma_commit(T_FULL(as));
                __syncwarp();
                if (++as == 2) { as = 0; t_phase ^= 1; }
            }
            if (leader) umma_commit(A_EMPTY);   // every MMA that reads this unit's query operand has retired
            __syncwarp();
        }
    } else if (warp >= 4 && warp < 4 + EPI_WARPS) {
        // ================================================================ epilogue
        const int ew = warp - 4, mt = ew >> 2, quarter = warp & 3;   // TMEM lanes 32*(warp%4) .. +31
        const int row = quarter * 32 + lane;                        // row inside the 128-query M tile
        const int gt = (ew & 3) * 32 + lane;                        // thread index inside the 128-thread group
        float *cn_buf = s_cn + mt * 2 * 128;
        const uint32_t bar_id = 1 + mt;
        int as = 0;
        uint32_t t_phase = 0;
        for (int u = blockIdx.x; u < p.nunits; u += gridDim.x) {
            const int qt = u / p.csplit, cs = u - qt * p.csplit;
            const int t0 = cs * p.tiles_per_unit;
            const int t1 = min(t0 + p.tiles_per_unit, p.ntile_c);
            const int q = qt * 128 * MT + mt * 128 + row;
            float f = q < p.nq ? p.qfac[q] : 0.f;
            // integer keys: everything is scaled by S (one scale per batch) and biased by the magic constant, so that
            // ONE fmaf per element yields a float whose bit pattern counts quanta of 128 / S
            float S = 1.f;
            if (IKEY) {
                S = coarse_key_scale(__uint_as_float(p.qstat[0]), __uint_as_float(p.cmax[1]));
                f *= S;
            }
            const float pad_cn = IKEY ? (COARSE_KEY_BIAS - 128.f + COARSE_KEY_MAGIC) : TC_INVALID;
            const uint32_t key_mul = p.key_mul;
            // ||c||^2 of a tile's column as staged in shared memory (integer keys: scaled and biased).  The global load
            // and the transform are kept apart: the load is issued a whole tile before its value is needed.
            auto load_cn = [&](int c) -> float { return c < p.nc ? p.cnorm[c] : TC_INVALID; };
            auto xform_cn = [&](float cn) -> float {
                if (!IKEY) return cn;
                return cn < 1.0e38f ? fmaf(cn, S, COARSE_KEY_MAGIC) : pad_cn;
            };
            // stage ||c||^2 of the first tile
            {
                cn_buf[gt] = xform_cn(load_cn(t0 * 128 + gt));
                asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
            }
            // TMEM -> register pipeline: ONE tcgen05.ld is always in flight.  Chunk jb of a tile lives in v[jb & 1];
            // while it is processed the next chunk -- or chunk 0 of the NEXT tile, whose accumulator the MMA warp
            // finished long ago -- is on its way.  (Left to itself ptxas sinks each load below the arithmetic of the
            // previous chunk -- it reuses the destination registers as temporaries -- and every chunk then pays the
            // full TMEM latency: 4 exposed round trips per tile were ~70 % of this kernel's time.  pin_regs() makes
            // the chunk's registers "written" after the next load has been issued, which fixes the order.)
            uint32_t v[2][32];
            auto pin_regs = [&](uint32_t (&r)[32]) {
                asm volatile("" : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                                  "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]),
                                  "+r"(r[15]), "+r"(r[16]), "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]),
                                  "+r"(r[22]), "+r"(r[23]), "+r"(r[24]), "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]),
                                  "+r"(r[29]), "+r"(r[30]), "+r"(r[31]) :: "memory");
            };
            auto tmem_addr = [&](int stage) -> uint32_t {
                return tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(stage * MT * 128 + mt * 128);
            };
            if (t0 < t1) {
                mbar_wait(T_FULL(as), t_phase);
                tc_fence_after();
                tmem_ld32(tmem_addr(as), v[0]);
            }
            // super-group record (gshift): the 4 smallest keys of up to 4 tiles, sub-tile in bits 7-8, value quantum 512
            uint32_t G0 = 0xffffffffu, G1 = 0xffffffffu, G2 = 0xffffffffu, G3 = 0xffffffffu;
            auto sg_insert = [&](uint32_t k, uint32_t sub) {
                uint32_t x = k >= COARSE_KEY_INVALID_MIN ? 0xffffffffu : (((k >> 9) << 9) | (sub << 7) | (k & 127u));
                const uint32_t m0 = max(G0, x); G0 = min(G0, x);
                const uint32_t m1 = max(G1, m0); G1 = min(G1, m0);
                const uint32_t m2 = max(G2, m1); G2 = min(G2, m1);
                G3 = min(G3, m2);
            };
            for (int t = t0; t < t1; t++) {
                const int b = (t - t0) & 1;
                const float cn_next = (t + 1 < t1) ? load_cn((t + 1) * 128 + gt) : TC_INVALID;
                const uint32_t taddr = tmem_addr(as);
                float a0 = TC_INVALID, a1 = TC_INVALID, a2 = TC_INVALID, a3 = TC_INVALID;   // KEEP == 3: the 4th slot stays empty
                uint32_t k0 = 0xffffffffu, k1 = 0xffffffffu, k2 = 0xffffffffu, k3 = 0xffffffffu;
#pragma unroll
                for (int jb = 0; jb < 4; jb++) {
                    tmem_ld_wait();
                    if (jb < 3) tmem_ld32(taddr + (jb + 1) * 32, v[(jb + 1) & 1]);
                    else {
                        // all TMEM reads of this accumulator stage are done: hand it back to the MMA warp ...
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(T_EMPTY(as));
                        // ... and start on the next tile's accumulator (the other stage)
                        if (t + 1 < t1) {
                            const int as2 = as ^ 1;
                            const uint32_t ph2 = as == 1 ? (t_phase ^ 1) : t_phase;
                            mbar_wait(T_FULL(as2), ph2);
                            tc_fence_after();
                            tmem_ld32(tmem_addr(as2), v[0]);
                        }
                    }
                    pin_regs(v[jb & 1]);
                    __syncwarp();     // a real instruction between the load and the arithmetic: ptxas sinks neither across it
                    const float4 *cn4 = reinterpret_cast<const float4 *>(cn_buf + b * 128 + jb * 32);
#pragma unroll
                    for (int i4 = 0; i4 < 8; i4++) {
                        const float4 cn = cn4[i4];
                        const float cnv[4] = {cn.x, cn.y, cn.z, cn.w};
                        uint32_t kprev = 0xffffffffu;
#pragma unroll
                        for (int ii = 0; ii < 4; ii++) {
                            const int i = i4 * 4 + ii;
                            if ((DIAG & 2) && i) continue;   // timing experiment: everything but the per-element work
                            const float val = fmaf(__uint_as_float(v[jb & 1][i]), f, cnv[ii]);
                            if (IKEY) {
                                // val in [2^30, 2^31): its bits count quanta.  (bits - bits(2^30)) * 128 + column in one
                                // multiply-add that wraps mod 2^32 (the FMA pipe; the ALU pipe keeps the min/max)
                                uint32_t key;
                                const uint32_t cst = (uint32_t)(jb * 32 + i) - (0x4e800000u << 7);
                                asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(key) : "r"(__float_as_uint(val)), "r"(key_mul), "r"(cst));
                                if (PRE == 2) {
                                    if (!(ii & 1)) { kprev = key; continue; }
                                    key = min(key, kprev);
                                }
                                const uint32_t m0 = max(k0, key);
                                k0 = min(k0, key);
                                if (KEEP == 2) k1 = min(k1, m0);
                                else {
                                    const uint32_t m1 = max(k1, m0);
                                    k1 = min(k1, m0);
                                    if (KEEP == 4) {
                                        const uint32_t m2 = max(k2, m1);
                                        k2 = min(k2, m1);
                                        k3 = min(k3, m2);
                                    } else k2 = min(k2, m1);
                                }
                            } else {
                                const float x = __uint_as_float((__float_as_uint(val) & p.keep_mask) | (uint32_t)(jb * 32 + i));
                                // sorted insert into the KEEP best of the group: 7 or 5 min/max
                                const float m0 = fmaxf(a0, x);
                                a0 = fminf(a0, x);
                                const float m1 = fmaxf(a1, m0);
                                a1 = fminf(a1, m0);
                                if (KEEP == 4) {
                                    const float m2 = fmaxf(a2, m1);
                                    a2 = fminf(a2, m1);
                                    a3 = fminf(a3, m2);
                                } else a2 = fminf(a2, m1);
                            }
                        }
                    }
                }
                if (IKEY && p.gshift) {
                    const uint32_t sub = (uint32_t)t & 3u;
                    sg_insert(k0, sub);
                    sg_insert(k1, sub);
                    if (KEEP >= 3) sg_insert(k2, sub);
                    if (KEEP >= 4) sg_insert(k3, sub);
                    if (sub == 3u || t == t1 - 1) {
                        if (q < p.nq)
                            *reinterpret_cast<uint4 *>(p.grp_val + ((size_t)q * ((p.ntile_c + 3) >> 2) + (t >> 2)) * COARSE_T) =
                                make_uint4(G0, G1, G2, G3);
                        G0 = G1 = G2 = G3 = 0xffffffffu;
                    }
                } else if (q < p.nq) {
                    if (IKEY)
                        *reinterpret_cast<uint4 *>(p.grp_val + ((size_t)q * p.ntile_c + t) * COARSE_T) =
                            make_uint4(k0, k1, k2, k3);
                    else
                        *reinterpret_cast<float4 *>(p.grp_val + ((size_t)q * p.ntile_c + t) * COARSE_T) =
                            make_float4(a0, a1, a2, a3);
                }
                cn_buf[(b ^ 1) * 128 + gt] = xform_cn(cn_next);
                asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
                if (++as == 2) { as = 0; t_phase ^= 1; }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
    }
}

// --------------------------------------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn()
{
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// rows x d fp16 matrix with row pitch dp, box = 64 columns x 128 rows, 128B swizzle, zero fill outside
static bool make_map(CUtensorMap *m, const void *base, size_t rows, int d, int dp)
{
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) return false;
    const cuuint64_t dims[2] = {(cuuint64_t)d, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)dp * 2};
    const cuuint32_t box[2] = {64, 128};
    const cuuint32_t estr[2] = {1, 1};
    return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void *>(base), dims, strides, box, estr,
              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

bool coarse_tc_supported(int d) { return d >= 16 && d <= 256; }

cudaError_t launch_coarse_tc(cudaStream_t st, uint64_t *launches, int nq, int nc, int d, const float *queries,
                             const float *cnorm_gemm, const TcSide &S, float *grp_val, int nterms, int keep, int ikey,
                             int gshift, int pre)
{
    if (!coarse_tc_supported(d)) return cudaErrorNotSupported;
    if (gshift && (!ikey || gshift != 2)) return cudaErrorInvalidValue;
    if (pre != 1 && (pre != 2 || !ikey || keep > 3)) return cudaErrorInvalidValue;
    // split the query rows (per-row power-of-two scale); batch statistics for the integer-key scale / select margin
    cudaError_t e = cudaMemsetAsync(S.qstat, 0, 8, st);
    if (e != cudaSuccess) return e;
    split_rows_kernel<<<(unsigned)((nq + 7) / 8), 256, 0, st>>>((size_t)nq, d, S.dp, queries, nullptr, S.cmax, S.qhi,
                                                               S.qlo, S.qfac, S.qstat, coarse_dims(d));
    ++*launches;
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;

    // only the dims fstdistfunc uses take part; the tensor maps zero-fill beyond them
    const int dk = coarse_dims(d);
    CUtensorMap m_qhi, m_qlo, m_chi, m_clo;
    if (!make_map(&m_qhi, S.qhi, nq, dk, S.dp) || !make_map(&m_qlo, S.qlo, nq, dk, S.dp) ||
        !make_map(&m_chi, S.chi, nc, dk, S.dp) || !make_map(&m_clo, S.clo, nc, dk, S.dp))
        return cudaErrorInvalidValue;

    TcParams p;
    p.nq = nq; p.nc = nc; p.d = dk; p.db = (dk + 63) / 64;
    p.nterms = nterms;
    p.qterms = ikey ? 1 : 2;
    p.cmax = S.cmax; p.qstat = S.qstat;
    const int MT = p.db <= 2 ? 2 : 1;
    p.ntile_c = (nc + 127) / 128;
    const int nqt = (nq + 128 * MT - 1) / (128 * MT);
    // pick the split of the centroid range that minimises the makespan over 148 persistent CTAs
    // (a unit costs its tiles plus ~1 tile for reloading the query operand)
    int best_cs = 1;
    double best = 1e30;
    const int max_cs = p.ntile_c < 256 ? p.ntile_c : 256;
    for (int cs = 1; cs <= max_cs; cs++) {
        const int tpu = (p.ntile_c + cs - 1) / cs;
        const int ncs = (p.ntile_c + tpu - 1) / tpu;
        const long units = (long)nqt * ncs;
        const double cost = (double)((units + 147) / 148) * (tpu + 1.0);
        if (cost < best - 1e-9) { best = cost; best_cs = ncs; }
    }
    p.tiles_per_unit = (p.ntile_c + best_cs - 1) / best_cs;
    if (gshift) p.tiles_per_unit = (p.tiles_per_unit + 3) & ~3;     // a super-group never straddles two units
    p.gshift = gshift;
    p.csplit = (p.ntile_c + p.tiles_per_unit - 1) / p.tiles_per_unit;
    p.nunits = nqt * p.csplit;
    const size_t a_bytes = (size_t)MT * p.db * 2 * TC_BLOCK_BYTES;
    const size_t fixed = 1024 + (size_t)MT * 2 * 128 * 4 + 32 * 8 + 16;
    int nst = (int)((232448 - fixed - a_bytes) / TC_BLOCK_BYTES);
    if (nst > 8) nst = 8;
    if (nst < 2) return cudaErrorNotSupported;
    p.nstages = nst;
    p.cnorm = cnorm_gemm; p.qfac = S.qfac; p.grp_val = grp_val; p.keep_mask = ~127u; p.key_mul = 128u;
    const size_t smem = fixed + a_bytes + (size_t)nst * TC_BLOCK_BYTES;
    const int grid = p.nunits < 148 ? p.nunits : 148;
#define IVF_TC_LAUNCH4(MTV, KV, IK, DG)                                                                                \
    do {                                                                                                               \
        e = cudaFuncSetAttribute(coarse_tc_kernel<MTV, KV, IK, DG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
        if (e != cudaSuccess) return e;                                                                                \
        coarse_tc_kernel<MTV, KV, IK, DG><<<grid, TC_THREADS, smem, st>>>(m_qhi, m_qlo, m_chi, m_clo, p);              \
    } while (0)
#define IVF_TC_LAUNCH5(MTV, KV, IK, DG, PR)                                                                            \
    do {                                                                                                               \
        e = cudaFuncSetAttribute(coarse_tc_kernel<MTV, KV, IK, DG, PR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
        if (e != cudaSuccess) return e;                                                                                \
        coarse_tc_kernel<MTV, KV, IK, DG, PR><<<grid, TC_THREADS, smem, st>>>(m_qhi, m_qlo, m_chi, m_clo, p);          \
    } while (0)
#define IVF_TC_LAUNCH(MTV, KV, IK)                                                                                     \
    do {                                                                                                               \
        e = cudaFuncSetAttribute(coarse_tc_kernel<MTV, KV, IK>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
        if (e != cudaSuccess) return e;                                                                                \
        coarse_tc_kernel<MTV, KV, IK><<<grid, TC_THREADS, smem, st>>>(m_qhi, m_qlo, m_chi, m_clo, p);                  \
    } while (0)
    static const int diag = getenv("IVFHNSW_B200_TC_DIAG") ? atoi(getenv("IVFHNSW_B200_TC_DIAG")) : 0;   // timing experiments
    if (diag && ikey && MT == 2 && keep == 2 && pre == 1) {
        if (diag == 1) IVF_TC_LAUNCH4(2, 2, true, 1);
        else if (diag == 2) IVF_TC_LAUNCH4(2, 2, true, 2);
        else IVF_TC_LAUNCH4(2, 2, true, 3);
    } else if (ikey && pre == 2) {
        if (MT == 2) { if (keep == 2) IVF_TC_LAUNCH5(2, 2, true, 0, 2); else IVF_TC_LAUNCH5(2, 3, true, 0, 2); }
        else { if (keep == 2) IVF_TC_LAUNCH5(1, 2, true, 0, 2); else IVF_TC_LAUNCH5(1, 3, true, 0, 2); }
    } else if (ikey) {
        if (MT == 2) { if (keep == 2) IVF_TC_LAUNCH(2, 2, true); else if (keep == 3) IVF_TC_LAUNCH(2, 3, true); else IVF_TC_LAUNCH(2, 4, true); }
        else { if (keep == 2) IVF_TC_LAUNCH(1, 2, true); else if (keep == 3) IVF_TC_LAUNCH(1, 3, true); else IVF_TC_LAUNCH(1, 4, true); }
    } else {
        if (MT == 2) { if (keep == 3) IVF_TC_LAUNCH(2, 3, false); else IVF_TC_LAUNCH(2, 4, false); }
        else { if (keep == 3) IVF_TC_LAUNCH(1, 3, false); else IVF_TC_LAUNCH(1, 4, false); }
    }
#undef IVF_TC_LAUNCH
#undef IVF_TC_LAUNCH4
#undef IVF_TC_LAUNCH5
    ++*launches;
    return cudaGetLastError();
}

}  // namespace ivfb200
