// ADC inverted-list scan for sm_100a: work-list construction (K-P / K-G1), the scan
// itself with the per-query lookup table held in shared memory (K-L1 + K-S1/K-S2) and
// the partial top-k merge (K-M1).
//
// Replaces, for a whole batch of queries at once, the body of
//   IndexIVF_HNSW::search            IndexIVF_HNSW.cpp:262-293   (and search2, :453-492)
//   IndexIVF_HNSW_Grouping::search   IndexIVF_HNSW_Grouping.cpp:222-353
//
// Data layout (what makes the table gathers bank-conflict free):
//   * device codes are stored per slot with their bytes XOR-permuted by the lane that will
//     read them: stored[t] = code[t ^ (slot & (min(M,32)-1))].  A warp reads 32 consecutive
//     slots, one per lane, with one 128-bit load each (M = 16).
//   * the per-query table lives in shared memory as 256 rows of 64 words (256 B).  Row j holds,
//     for every sub-quantizer m, the value -2<q_m, c_{m,j}> at word R*m + copy (R = 32/M copies)
//     and in words 32..63 the decoded norm norm_table[j] replicated 32 times.
//     At step t lane l needs sub-quantizer m = t ^ (l % M), which it finds at word
//     R*m + l/M: for a fixed t these 32 words are distinct banks whatever the code bytes are,
//     and the norm lookup (word 32 + l) is conflict free as well.
//   * a lookup address is built with ONE prmt: byte1 = code byte, byte0 = per-lane word offset.
//
// A CTA (8 warps) owns one (query, split); warps pull work items -- (list, part of <= SCAN_SEG positions) --
// from a shared counter.  An item is processed as 16-wide WINDOWS that may start at any list position (two per
// warp step, one per half warp; 32-wide for M >= 32): with pruning only the runs of kept sub-groups are tiled.
// Candidates pass a float prefilter, go to a per-warp buffer and are merged into the warp's register-resident
// sorted list a batch at a time (k <= 32), inserted one by one (k <= 128) or kept in a locked shared list.
#include "common.cuh"
#include "kernels.h"
#include <type_traits>

namespace ivfb200 {

// =============================================================================================
// K-P: IVFADC work list.  One warp per query.  Reproduces the sequential max_codes rule of
// IndexIVF_HNSW.cpp:267-293: lists are taken in rank order, empty lists are skipped, and the loop
// stops AFTER the list that makes ncode >= max_codes  <=>  list i is scanned iff the sizes of the
// lists before it sum to < max_codes.
// =============================================================================================
__global__ void prep_adc_kernel(int nq, int nprobe, uint64_t max_codes, const float *__restrict__ coarse_d,
                                const uint32_t *__restrict__ coarse_i, DeviceLists L, SegTables T)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = blockIdx.x * (blockDim.x >> 5) + warp;
    if (q >= nq) return;
    uint64_t run = 0;
    uint32_t scanned = 0;
    for (int base = 0; base < nprobe; base += 32) {
        const int i = base + lane;
        uint32_t c = 0, sz = 0;
        float cd = 0.f;
        if (i < nprobe) {
            c = coarse_i[(size_t)q * nprobe + i];
            cd = coarse_d[(size_t)q * nprobe + i];
            sz = L.list_size[c];
        }
        uint64_t incl = sz;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const uint64_t t = __shfl_up_sync(IVF_FULL_MASK, incl, off);
            if (lane >= off) incl += t;
        }
        const uint64_t excl = run + incl - sz;
        const bool active = (i < nprobe) && sz > 0 && excl < max_codes;
        const uint32_t n = active ? sz : 0u;
        if (i < nprobe) {
            T.slot(q)[i] = L.list_slot[c];
            T.n(q)[i] = n;
            T.base(q)[i] = cd - L.centroid_norms[c];
            T.seq(q)[i] = (uint32_t)excl;
            T.cid(q)[i] = c;
        }
        run += __shfl_sync(IVF_FULL_MASK, incl, 31);
        uint32_t s = n;
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) s += __shfl_xor_sync(IVF_FULL_MASK, s, off);
        scanned += s;
    }
    if (lane == 0) {
        T.nscanned[q] = scanned;
        T.nwalked[q] = scanned;
        T.threshold[q] = 0.f;
    }
}

cudaError_t launch_prep_adc(cudaStream_t st, uint64_t *launches, int nq, int nprobe, uint64_t max_codes,
                            const float *coarse_d, const uint32_t *coarse_i, const DeviceLists &L, const SegTables &T)
{
    if (nq == 0) return cudaSuccess;
    const int wpb = 4;
    prep_adc_kernel<<<(nq + wpb - 1) / wpb, wpb * 32, 0, st>>>(nq, nprobe, max_codes, coarse_d, coarse_i, L, T);
    ++*launches;
    return cudaGetLastError();
}

// =============================================================================================
// Grouping work list (IndexIVF_HNSW_Grouping.cpp:222-262 and :283-353), three kernels:
//
//  K-G1  distances from the query to the neighbour centroids n_s of every probed list (the lazily memoised
//        ivfhnsw::fvec_L2sqr calls of :246-250 and :312-316), three kernels:
//        K-G1a one CTA per query.  Like the reference's nc-sized memo (query_centroid_dists, :57) every
//              distinct centroid is evaluated ONCE per query: the neighbour ids of all probed lists are
//              de-duplicated in a shared-memory hash table (the 64-NN lists of the probed centroids overlap
//              heavily) into a dense row list.  Memo semantics kept: a neighbour that is itself a probed
//              centroid reuses the coarse distance (:207-220) unless that value is < EPS.
//        K-G1b the distances: whole rows fetched with coalesced cp.async into a shared-memory ring, two
//              lanes per row adding the terms in fvec_L2sqr's order.
//        K-G1c per (probe, sub-group) pair:
//        Out: T.qsd[q][i][s]      = the pruning value query_subcentroid_dists (:251-252), NaN where it
//                                   does not contribute to the threshold
//             T.sub_base[q][i][s] = term1 + term2 of the sub-group (+inf for empty ones)
//             T.nscanned[q]       = number of contributing values (scratch for K-G2)
//  K-G2  pruning threshold = SEQUENTIAL fp32 sum of the contributing values in (probe, sub-group) order
//        divided by their count.  The chain is inherently serial per query, so a CTA
//        runs 32 queries side by side: tiles are transposed through shared memory and lane <-> query.
//  K-G3  one CTA per query: applies the threshold (rows past the 2*max_codes horizon compare 0, :228),
//        counts the kept codes per list, applies the max_codes rule on the KEPT codes (:337,349) and
//        writes the segment tables.
// =============================================================================================
#define G1_THREADS 512

struct ProbeMeta {
    uint32_t *c;      // centroid id
    uint32_t *sz;     // list size
    uint8_t *inh;     // inside the pass-1 horizon
};

// pass-1 horizon (:229-259): list i is visited iff non-empty and the sizes before it sum to < 2*max_codes.
// Whole CTA: every probe's (centroid id, list size) is fetched by its own thread -- ONE round of dependent global loads
// for all probes -- then warp 0 runs the prefix over shared memory.  (One warp doing both, 32 probes per trip, left
// the other 15 warps of the three preparation kernels at the barrier for 4 serial round trips: 25-35 % of K-G1c/K-G3.)
// Contains a __syncthreads(); the caller syncs again before reading the results.
struct ProbeNoExtra {
    __device__ __forceinline__ void operator()(int, uint32_t) const {}
};
template <class Extra = ProbeNoExtra>   // extra(i, c): further per-probe loads that ride the same round trip
static __device__ __forceinline__ void probe_meta_cta(int nprobe, uint64_t max_codes, const uint32_t *coarse_i_q,
                                                      const DeviceLists &L, uint32_t *s_c, uint32_t *s_sz,
                                                      uint8_t *s_inh, int tid, int nthreads, Extra extra = Extra())
{
    for (int i = tid; i < nprobe; i += nthreads) {
        const uint32_t c = coarse_i_q[i];
        const uint32_t sz = L.list_size[c];
        extra(i, c);
        s_c[i] = c;
        s_sz[i] = sz;
    }
    __syncthreads();
    if (tid >= 32) return;
    const int lane = tid;
    uint64_t run = 0;
    for (int base = 0; base < nprobe; base += 32) {
        const int i = base + lane;
        const uint32_t sz = i < nprobe ? s_sz[i] : 0u;
        uint64_t incl = sz;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const uint64_t t = __shfl_up_sync(IVF_FULL_MASK, incl, off);
            if (lane >= off) incl += t;
        }
        const uint64_t excl = run + incl - sz;
        if (i < nprobe) s_inh[i] = (sz > 0 && excl < 2 * max_codes) ? 1 : 0;
        run += __shfl_sync(IVF_FULL_MASK, incl, 31);
    }
}

static __device__ __forceinline__ void ldg_256(const float *p, float (&v)[8])
{
    asm volatile("ld.global.nc.v8.f32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
                 : "l"(p));
}

// ivfhnsw::fvec_L2sqr (utils.cpp:22-51) by ONE thread: 8 partial sums over floor(d/16)*16 dims, unfused
// multiply-add like the pinned C restatement, partials added left to right.  x in shared memory.
static __device__ __forceinline__ float l2sqr16_thread(const float *__restrict__ x, const float *__restrict__ y, int d)
{
    const int n = (d >> 4) << 4;
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; j++) acc[j] = 0.f;
    if ((d & 7) == 0) {
#pragma unroll 4
        for (int i = 0; i < n; i += 8) {
            float yv[8];
            ldg_256(y + i, yv);
            const float4 x0 = *reinterpret_cast<const float4 *>(x + i), x1 = *reinterpret_cast<const float4 *>(x + i + 4);
            const float xv[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const float df = xv[j] - yv[j];
                acc[j] = __fadd_rn(acc[j], __fmul_rn(df, df));
            }
        }
    } else {
        for (int i = 0; i < n; i += 8) {
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const float df = x[i + j] - __ldg(y + i + j);
                acc[j] = __fadd_rn(acc[j], __fmul_rn(df, df));
            }
        }
    }
    float r = acc[0];
#pragma unroll
    for (int j = 1; j < 8; j++) r += acc[j];
    return r;
}

// e -> (probe, sub-group) without an integer division when nsubc is a power of two (the usual 64)
static __device__ __forceinline__ int pair_probe(int e, int nsubc, int sub_shift)
{
    return sub_shift >= 0 ? (e >> sub_shift) : (e / nsubc);
}

// ---- K-G1a: distinct neighbour centroids of a query (hash de-duplication) -> dense row list + pair map
__global__ void __launch_bounds__(G1_THREADS) group_g1a_kernel(int nprobe, int nsubc, int sub_shift, int ht_size,
                                                               uint64_t max_codes,
                                                               const float *__restrict__ coarse_d,
                                                               const uint32_t *__restrict__ coarse_i, DeviceLists L,
                                                               G1Work Wk, unsigned long long *rows_counter)
{
    extern __shared__ __align__(16) uint8_t sm[];
    uint32_t *hk = reinterpret_cast<uint32_t *>(sm);                   // [ht_size] centroid ids
    uint32_t *s_c = hk + ht_size;                                      // [nprobe]
    uint32_t *s_sz = s_c + nprobe;
    uint16_t *hidx = reinterpret_cast<uint16_t *>(s_sz + nprobe);      // [ht_size] dense row index of the slot
    uint16_t *pslot = hidx + ht_size;                                  // [npairs] table slot of every pair
    uint8_t *s_inh = reinterpret_cast<uint8_t *>(pslot + (size_t)nprobe * nsubc);
    __shared__ int s_nwork, s_nmemo;
    const int q = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const size_t qo = (size_t)q * nprobe;
    const uint32_t H = (uint32_t)ht_size;
    uint32_t *rows = Wk.rows + (size_t)q * Wk.cap;
    float2 *dv = Wk.dv + (size_t)q * Wk.cap;
    for (int i = tid; i < ht_size; i += G1_THREADS) hk[i] = 0xffffffffu;
    if (tid == 0) { s_nwork = 0; s_nmemo = 0; }
    probe_meta_cta(nprobe, max_codes, coarse_i + qo, L, s_c, s_sz, s_inh, tid, G1_THREADS);
    __syncthreads();
    // ---- the coarse stage already knows d(q, c) of the probed centroids: they enter the table as done (the
    // reference's query_centroid_dists memo), unless the value is < EPS, which the reference treats as "not
    // computed" (Q1 iv).  Of a centroid listed twice the last entry counts (it overwrites).
    for (int i = tid; i < nprobe; i += G1_THREADS) {
        const uint32_t c = s_c[i];
        const float v = coarse_d[qo + i];
        bool last = true;
        for (int j = i + 1; j < nprobe; j++) last = last && (s_c[j] != c);
        if (last && v >= 0.00001f) {
            uint32_t p = __umulhi(c * 2654435761u, H);
            while (atomicCAS(&hk[p], 0xffffffffu, c) != 0xffffffffu) p = p + 1 == H ? 0u : p + 1;
            const int m = atomicAdd(&s_nmemo, 1);
            hidx[p] = (uint16_t)m;
            rows[m] = c;
            dv[m] = make_float2(v, L.centroid_norms[c]);
        }
    }
    __syncthreads();
    const int npairs = nprobe * nsubc, nmemo = s_nmemo;
    // ---- every needed neighbour id enters the table once (loads of 8 pairs in flight per thread); a pair
    // remembers its slot, first sightings get the next dense index (one counter update per warp)
    for (int e0 = 0; e0 < npairs; e0 += G1_THREADS * 8) {
        uint32_t nn[8];
#pragma unroll
        for (int u = 0; u < 8; u++) {
            const int e = e0 + u * G1_THREADS + tid;
            nn[u] = 0xffffffffu;
            if (e < npairs) {
                const int i = pair_probe(e, nsubc, sub_shift), s = e - i * nsubc;
                if (s_sz[i] != 0) {
                    const size_t o = (size_t)s_c[i] * nsubc + s;
                    const uint32_t ss = __ldg(L.sub_sizes + o);
                    const uint32_t n = __ldg(L.nn_idx + o);
                    if (ss != 0) nn[u] = n;
                }
            }
        }
#pragma unroll
        for (int u = 0; u < 8; u++) {
            const int e = e0 + u * G1_THREADS + tid;
            const uint32_t n = nn[u];
            uint32_t p = 0xffffu;
            bool first = false;
            if (n != 0xffffffffu) {
                p = __umulhi(n * 2654435761u, H);
                while (true) {
                    const uint32_t prev = atomicCAS(&hk[p], 0xffffffffu, n);
                    if (prev == 0xffffffffu) { first = true; break; }
                    if (prev == n) break;
                    p = p + 1 == H ? 0u : p + 1;
                }
            }
            const uint32_t fm = __ballot_sync(IVF_FULL_MASK, first);
            if (fm) {
                const int leader = __ffs(fm) - 1;
                int base = 0;
                if (lane == leader) base = atomicAdd(&s_nwork, __popc(fm));
                base = __shfl_sync(IVF_FULL_MASK, base, leader);
                if (first) {
                    const int w = nmemo + base + __popc(fm & ((1u << lane) - 1u));
                    hidx[p] = (uint16_t)w;
                    rows[w] = n;
                }
            }
            if (e < npairs) pslot[e] = (uint16_t)p;
        }
    }
    __syncthreads();
    uint16_t *ppos = Wk.ppos + (size_t)q * npairs;
    for (int e = tid; e < npairs; e += G1_THREADS) {
        const uint16_t p = pslot[e];
        ppos[e] = p == 0xffffu ? (uint16_t)0xffffu : hidx[p];
    }
    if (tid == 0) {
        Wk.cnt[q] = make_uint2((uint32_t)nmemo, (uint32_t)s_nwork);
        if (rows_counter) atomicAdd(rows_counter, (unsigned long long)s_nwork);
    }
}

// ---- K-G1b: d(q, row) of the rows K-G1a listed.  The rows live in L2 (or HBM): what matters is the shape of the
// requests and the bytes in flight.  A warp fetches whole rows with cp.async (16 bytes per lane: full 128-byte
// lines, 3x the rows/s of per-thread sector loads -- profiles/scratch/gather_bench2.cu) into a two-stage ring of
// 16-row tiles in shared memory; while one tile is in flight the other is summed, TWO lanes per row: lane 2r
// owns partial sums 0..3 of row r and lane 2r+1 partial sums 4..7, each added in EXACTLY the order of
// ivfhnsw::fvec_L2sqr (utils.cpp:22-51) -- 8 partial sums over floor(d/16)*16 dims, unfused multiply-add like the
// pinned C restatement -- and the partials are then added left to right.  Row pitch n + 8 words = 8 * odd:
// the 128-bit reads of a quarter warp fall in 8 distinct bank groups.
#define G1B_WARPS 4
#define G1B_TILE 16
static __device__ __forceinline__ void cp_async16(void *smem_dst, const void *gsrc)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc)
                 : "memory");
}
template <int N>   // dims fvec_L2sqr uses, compile time (query in registers); 0 = run time (query in shared memory)
__global__ void __launch_bounds__(G1B_WARPS * 32, 3) group_g1b_kernel(int d, int n_rt, const float *__restrict__ queries,
                                                                      DeviceLists L, G1Work Wk)
{
    extern __shared__ __align__(16) float g1b_sm[];
    const int n = N > 0 ? N : n_rt;
    const int q = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint2 cnt = Wk.cnt[q];
    const int nmemo = (int)cnt.x, nwork = (int)cnt.y;
    const int pitch = n + 8;
    float *buf = g1b_sm + (size_t)warp * 2 * G1B_TILE * pitch;
    float *xs = g1b_sm + (size_t)G1B_WARPS * 2 * G1B_TILE * pitch;     // N == 0: the query
    const uint32_t *rows = Wk.rows + (size_t)q * Wk.cap + nmemo;
    float2 *dv = Wk.dv + (size_t)q * Wk.cap + nmemo;
    const int h = lane & 1, r = lane >> 1;
    float4 xr[N > 0 ? N / 8 : 1];
    if (N > 0) {
#pragma unroll
        for (int i = 0; i < N / 8; i++) xr[i] = *reinterpret_cast<const float4 *>(queries + (size_t)q * d + 8 * i + 4 * h);
    } else {
        for (int i = threadIdx.x; i < n; i += G1B_WARPS * 32) xs[i] = queries[(size_t)q * d + i];
        __syncthreads();
    }
    const int ntiles = (nwork + G1B_TILE - 1) / G1B_TILE;
    // tile t of this warp: rows [t * TILE, ...); lane l copies bytes [16 l, 16 l + 16) (+512 per pass) of every row
    auto issue = [&](int t, int stage) {
        const int b = t * G1B_TILE;
        const int nb = min(G1B_TILE, nwork - b);
        const uint32_t my = lane < nb ? rows[b + lane] : 0u;
        float *dst = buf + (size_t)stage * G1B_TILE * pitch;
        for (int u = 0; u < nb; u++) {
            const uint32_t row = __shfl_sync(IVF_FULL_MASK, my, u);
            const float *src = L.centroids + (size_t)row * d;
            for (int o = 4 * lane; o < n; o += 128) cp_async16(dst + (size_t)u * pitch + o, src + o);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    int t = warp;
    if (t < ntiles) issue(t, 0);
    for (int it = 0; t < ntiles; t += G1B_WARPS, it++) {
        const int stage = it & 1;
        if (t + G1B_WARPS < ntiles) {
            issue(t + G1B_WARPS, stage ^ 1);
            asm volatile("cp.async.wait_group 1;" ::: "memory");
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        __syncwarp();
        const int b = t * G1B_TILE;
        const int nb = min(G1B_TILE, nwork - b);
        const float *pr = buf + ((size_t)stage * G1B_TILE + r) * pitch + 4 * h;
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
        if (N > 0) {
#pragma unroll
            for (int i = 0; i < N / 8; i++) {
                const float4 y = *reinterpret_cast<const float4 *>(pr + 8 * i);
                const float d0 = xr[i].x - y.x, d1 = xr[i].y - y.y, d2 = xr[i].z - y.z, d3 = xr[i].w - y.w;
                a0 = __fadd_rn(a0, __fmul_rn(d0, d0));
                a1 = __fadd_rn(a1, __fmul_rn(d1, d1));
                a2 = __fadd_rn(a2, __fmul_rn(d2, d2));
                a3 = __fadd_rn(a3, __fmul_rn(d3, d3));
            }
        } else {
#pragma unroll 4
            for (int i = 0; i < n; i += 8) {
                const float4 y = *reinterpret_cast<const float4 *>(pr + i);
                const float4 x = *reinterpret_cast<const float4 *>(xs + i + 4 * h);
                const float d0 = x.x - y.x, d1 = x.y - y.y, d2 = x.z - y.z, d3 = x.w - y.w;
                a0 = __fadd_rn(a0, __fmul_rn(d0, d0));
                a1 = __fadd_rn(a1, __fmul_rn(d1, d1));
                a2 = __fadd_rn(a2, __fmul_rn(d2, d2));
                a3 = __fadd_rn(a3, __fmul_rn(d3, d3));
            }
        }
        // acc[0] + acc[1] + ... + acc[7], left to right: the even lane's four, then the odd lane's
        const float lo4 = ((a0 + a1) + a2) + a3;
        const float left = __shfl_sync(IVF_FULL_MASK, lo4, lane & ~1);
        const float res = (((left + a0) + a1) + a2) + a3;
        if (h == 1 && r < nb) {
            const uint32_t row = rows[b + r];
            dv[b + r] = make_float2(res, __ldg(L.centroid_norms + row));
        }
        __syncwarp();
    }
}

// shapes the transposing kernel does not take (d % 4 != 0 or very long rows): one thread per row
__global__ void __launch_bounds__(256) group_g1b_simple_kernel(int d, const float *__restrict__ queries, DeviceLists L,
                                                               G1Work Wk)
{
    extern __shared__ __align__(16) float g1b_sq[];
    const int q = blockIdx.x;
    const uint2 cnt = Wk.cnt[q];
    for (int i = threadIdx.x; i < d; i += blockDim.x) g1b_sq[i] = queries[(size_t)q * d + i];
    __syncthreads();
    const uint32_t *rows = Wk.rows + (size_t)q * Wk.cap + cnt.x;
    float2 *dv = Wk.dv + (size_t)q * Wk.cap + cnt.x;
    for (int w = threadIdx.x; w < (int)cnt.y; w += blockDim.x) {
        const uint32_t n = rows[w];
        dv[w] = make_float2(l2sqr16_thread(g1b_sq, L.centroids + (size_t)n * d, d), __ldg(L.centroid_norms + n));
    }
}

// ---- K-G1c: per sub-group pruning value and base (G1C_U pairs per thread in flight: the pair-map load, then the
// dependent (dist, norm) gather, bound this kernel by latency)
#define G1C_U 8
__global__ void __launch_bounds__(G1_THREADS) group_g1c_kernel(int nprobe, int nsubc, int sub_shift, uint64_t max_codes,
                                                               int do_pruning,
                                                               const float *__restrict__ coarse_d,
                                                               const uint32_t *__restrict__ coarse_i, DeviceLists L,
                                                               SegTables T, G1Work Wk)
{
    extern __shared__ __align__(16) uint8_t sm[];
    uint32_t *s_c = reinterpret_cast<uint32_t *>(sm);      // [nprobe]
    uint32_t *s_sz = s_c + nprobe;
    float *s_alpha = reinterpret_cast<float *>(s_sz + nprobe);
    float *s_cd = s_alpha + nprobe;
    float *s_cn = s_cd + nprobe;
    uint8_t *s_inh = reinterpret_cast<uint8_t *>(s_cn + nprobe);
    __shared__ uint32_t s_count;
    const int q = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const size_t qo = (size_t)q * nprobe;
    const float qnan = __int_as_float(0x7fc00000);
    if (tid == 0) s_count = 0;
    probe_meta_cta(nprobe, max_codes, coarse_i + qo, L, s_c, s_sz, s_inh, tid, G1_THREADS, [&](int i, uint32_t c) {
        const float al = L.alphas[c], cd = coarse_d[qo + i], cn = L.centroid_norms[c];
        s_alpha[i] = al;
        s_cd[i] = cd;
        s_cn[i] = cn;
    });
    __syncthreads();
    const int npairs = nprobe * nsubc;
    const uint16_t *ppos = Wk.ppos + (size_t)q * npairs;
    // (dist, norm) pairs are gathered straight from global memory: staging them in shared memory was measured
    // slower (occupancy, bank conflicts: C2 0.25 -> 0.39 ms)
    const float2 *__restrict__ dv = Wk.dv + (size_t)q * Wk.cap;
    uint32_t count = 0;
    for (int e0 = 0; e0 < npairs; e0 += G1_THREADS * G1C_U) {
        uint32_t pos[G1C_U];
        float icd[G1C_U];
        float2 v[G1C_U];
#pragma unroll
        for (int u = 0; u < G1C_U; u++) {
            const int e = e0 + u * G1_THREADS + tid;
            pos[u] = e < npairs ? ppos[e] : 0xffffu;
        }
#pragma unroll
        for (int u = 0; u < G1C_U; u++) {
            const int e = e0 + u * G1_THREADS + tid;
            icd[u] = 0.f;
            v[u] = make_float2(0.f, 0.f);
            if (pos[u] != 0xffffu) {
                const int i = pair_probe(e, nsubc, sub_shift), s = e - i * nsubc;
                icd[u] = __ldg(L.icd + (size_t)s_c[i] * nsubc + s);
                v[u] = dv[pos[u]];
            }
        }
#pragma unroll
        for (int u = 0; u < G1C_U; u++) {
            const int e = e0 + u * G1_THREADS + tid;
            if (e >= npairs) continue;
            float r = qnan, sb = __int_as_float(0x7f800000);
            if (pos[u] != 0xffffu) {
                const int i = pair_probe(e, nsubc, sub_shift);
                const float qsd = v[u].x;
                const float cnn = v[u].y;
                const float alpha = s_alpha[i];
                const float oma = __fsub_rn(1.f, alpha);
                // :251-252  (1-a)*d(q,c) - a*((1-a)*d(c,n) - d(q,n))
                if (do_pruning && s_inh[i]) {
                    const float t1 = __fmul_rn(oma, s_cd[i]);
                    const float inner = __fsub_rn(__fmul_rn(oma, icd[u]), qsd);
                    r = __fsub_rn(t1, __fmul_rn(alpha, inner));
                    count++;
                }
                // :318-322  term1 + term2
                const float term1 = __fmul_rn(oma, __fsub_rn(s_cd[i], s_cn[i]));
                const float term2 = __fmul_rn(alpha, __fsub_rn(qsd, cnn));
                sb = __fadd_rn(term1, term2);
            }
            T.qsd[(size_t)q * npairs + e] = r;
            T.sub_base(q)[e] = sb;
        }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) count += __shfl_xor_sync(IVF_FULL_MASK, count, off);
    if (lane == 0 && count) atomicAdd(&s_count, count);
    __syncthreads();
    if (tid == 0) T.nscanned[q] = s_count;
}

// K-G2: lane <-> query, 32 queries per CTA, tiles of G2_TE values per query transposed through smem
#define G2_TE 64
__global__ void __launch_bounds__(128) group_g2_kernel(int nq, int npairs, const float *__restrict__ vals,
                                                       const uint32_t *__restrict__ counts, float *__restrict__ thr)
{
    __shared__ float tile[2][32][G2_TE + 1];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int q0 = blockIdx.x * 32;
    const int ntiles = (npairs + G2_TE - 1) / G2_TE;
    const float qnan = __int_as_float(0x7fc00000);
    // thread t loads element (row = (t*? ) ...): 32 rows x 64 columns = 2048 values, 16 per thread,
    // row r = idx / 64, col = idx % 64 with idx = u * 128 + tid  -> a warp reads 32 consecutive columns
    float reg[16];
    auto load_tile = [&](int t) {
#pragma unroll
        for (int u = 0; u < 16; u++) {
            const int idx = u * 128 + tid, r = idx >> 6, c = idx & 63;
            const int e = t * G2_TE + c, q = q0 + r;
            reg[u] = (q < nq && e < npairs) ? __ldg(vals + (size_t)q * npairs + e) : qnan;
        }
    };
    float acc = 0.f;
    load_tile(0);
    for (int t = 0; t < ntiles; t++) {
        const int b = t & 1;
#pragma unroll
        for (int u = 0; u < 16; u++) {
            const int idx = u * 128 + tid;
            tile[b][idx >> 6][idx & 63] = reg[u];
        }
        __syncthreads();
        if (t + 1 < ntiles) load_tile(t + 1);
        if (warp == 0) {
#pragma unroll 16
            for (int c = 0; c < G2_TE; c++) {
                const float v = tile[b][lane][c];
                if (v == v) acc = __fadd_rn(acc, v);
            }
        }
    }
    if (warp == 0 && q0 + lane < nq) thr[q0 + lane] = acc / (float)counts[q0 + lane];   // 0/0 -> NaN like the reference
}

__global__ void __launch_bounds__(G1_THREADS) group_g3_kernel(int nprobe, int nsubc, int sub_shift, uint64_t max_codes,
                                                              int do_pruning,
                                                              const uint32_t *__restrict__ coarse_i, DeviceLists L,
                                                              SegTables T)
{
    extern __shared__ __align__(16) uint8_t sm[];
    uint32_t *s_c = reinterpret_cast<uint32_t *>(sm);     // [nprobe] centroid id
    uint32_t *s_sz = s_c + nprobe;                        // [nprobe] list size
    uint32_t *s_kept = s_sz + nprobe;                     // [nprobe] kept codes
    uint8_t *s_inh = reinterpret_cast<uint8_t *>(s_kept + nprobe);
    const int q = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const size_t qo = (size_t)q * nprobe;
    const int npairs = nprobe * nsubc;
    probe_meta_cta(nprobe, max_codes, coarse_i + qo, L, s_c, s_sz, s_inh, tid, G1_THREADS);
    for (int i = tid; i < nprobe; i += G1_THREADS) s_kept[i] = 0;
    __syncthreads();
    const float threshold = do_pruning ? T.threshold[q] : 0.f;
    const float *rv = T.qsd + (size_t)q * npairs;
    float *sb = T.sub_base(q);
    for (int e0 = 0; e0 < npairs; e0 += G1_THREADS * 4) {
        uint32_t ss[4];
        float r[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int e = e0 + u * G1_THREADS + tid;
            ss[u] = 0;
            r[u] = 0.f;
            if (e < npairs) {
                const int i = pair_probe(e, nsubc, sub_shift), s = e - i * nsubc;
                if (s_sz[i] != 0) {
                    ss[u] = __ldg(L.sub_sizes + (size_t)s_c[i] * nsubc + s);
                    if (do_pruning && s_inh[i]) r[u] = rv[e];    // rows past the horizon stay 0 (:228)
                }
            }
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int e = e0 + u * G1_THREADS + tid;
            if (e >= npairs || ss[u] == 0) continue;
            const bool keep = !do_pruning || r[u] < threshold;
            if (keep) atomicAdd(&s_kept[pair_probe(e, nsubc, sub_shift)], ss[u]);
            else sb[e] = __int_as_float(0x7f800000);
        }
    }
    __syncthreads();
    if (warp != 0) return;
    // ---- max_codes rule on the KEPT codes (:337,349) and segment tables
    uint64_t run_k = 0, run_w = 0;
    uint32_t scanned = 0, walked = 0;
    for (int base = 0; base < nprobe; base += 32) {
        const int i = base + lane;
        uint32_t c = 0, sz = 0, kp = 0;
        if (i < nprobe) { c = s_c[i]; sz = s_sz[i]; kp = s_kept[i]; }
        uint64_t incl_k = kp;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const uint64_t t = __shfl_up_sync(IVF_FULL_MASK, incl_k, off);
            if (lane >= off) incl_k += t;
        }
        const uint64_t excl_k = run_k + incl_k - kp;
        const bool active = (i < nprobe) && sz > 0 && excl_k < max_codes;
        const uint32_t wsz = active ? sz : 0u;  // walk order numbering covers every active list
        uint64_t incl_w = wsz;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const uint64_t t = __shfl_up_sync(IVF_FULL_MASK, incl_w, off);
            if (lane >= off) incl_w += t;
        }
        const uint32_t n = (active && kp > 0) ? sz : 0u;
        if (i < nprobe) {
            T.slot(q)[i] = L.list_slot[c];
            T.n(q)[i] = n;
            T.base(q)[i] = 0.f;
            T.seq(q)[i] = (uint32_t)(run_w + incl_w - wsz);
            T.cid(q)[i] = c;
        }
        run_k += __shfl_sync(IVF_FULL_MASK, incl_k, 31);
        run_w += __shfl_sync(IVF_FULL_MASK, incl_w, 31);
        uint32_t a = active ? kp : 0u, b = n;
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            a += __shfl_xor_sync(IVF_FULL_MASK, a, off);
            b += __shfl_xor_sync(IVF_FULL_MASK, b, off);
        }
        scanned += a;
        walked += b;
    }
    if (lane == 0) {
        T.nscanned[q] = scanned;
        T.nwalked[q] = walked;
        if (!do_pruning) T.threshold[q] = 0.f;
    }
}

size_t g1_work_bytes(int nprobe, int nsubc)
{
    const size_t cap = (size_t)nprobe * nsubc + nprobe;
    return cap * 12 + (size_t)nprobe * nsubc * 2 + 16;      // rows + (dist, norm) + pair map + counts, per query
}

template <int N>
static cudaError_t launch_g1b(cudaStream_t st, int nq, int d, int n, const float *queries, const DeviceLists &L,
                              const G1Work &Wk)
{
    const size_t smb = (size_t)G1B_WARPS * 2 * G1B_TILE * (n + 8) * 4 + (N > 0 ? 0 : (size_t)n * 4);
    cudaError_t e = cudaFuncSetAttribute(group_g1b_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smb);
    if (e != cudaSuccess) return e;
    group_g1b_kernel<N><<<nq, G1B_WARPS * 32, smb, st>>>(d, n, queries, L, Wk);
    return cudaGetLastError();
}

cudaError_t launch_prep_grouping(cudaStream_t st, uint64_t *launches, int nq, int d, int nprobe, int nsubc,
                                 uint64_t max_codes, bool do_pruning, const float *queries,
                                 const float *coarse_d, const uint32_t *coarse_i, const DeviceLists &L,
                                 const SegTables &T, void *g1_work, cudaEvent_t ev_before_g1b, cudaEvent_t ev_after_g1,
                                 unsigned long long *rows_counter)
{
    if (nq == 0) return cudaSuccess;
    const int npairs = nprobe * nsubc;
    // table: the memoised probes + one slot per (probe, sub-group) pair at most + one that stays empty
    // (typical load 20-30 %: the neighbour lists of the probed centroids overlap)
    const int ht = npairs + nprobe + 1;
    if (ht > 65535) return cudaErrorInvalidValue;       // uint16 slot numbers
    G1Work Wk;
    Wk.cap = npairs + nprobe;
    {   // carve the per-batch workspace: [cnt][dv][rows][ppos]
        uint8_t *p = static_cast<uint8_t *>(g1_work);
        Wk.cnt = reinterpret_cast<uint2 *>(p); p += (size_t)nq * 8;
        Wk.dv = reinterpret_cast<float2 *>(p); p += (size_t)nq * Wk.cap * 8;
        Wk.rows = reinterpret_cast<uint32_t *>(p); p += (size_t)nq * Wk.cap * 4;
        Wk.ppos = reinterpret_cast<uint16_t *>(p);
    }
    int sub_shift = -1;
    for (int b = 0; b < 16; b++)
        if ((1 << b) == nsubc) sub_shift = b;
    const size_t sm1 = (size_t)ht * 6 + (size_t)nprobe * 8 + (size_t)npairs * 2 + (size_t)nprobe + 16;
    cudaError_t e = cudaFuncSetAttribute(group_g1a_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm1);
    if (e != cudaSuccess) return e;
    group_g1a_kernel<<<nq, G1_THREADS, sm1, st>>>(nprobe, nsubc, sub_shift, ht, max_codes, coarse_d, coarse_i, L, Wk,
                                                 rows_counter);
    ++*launches;
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    if (ev_before_g1b) {
        e = cudaEventRecord(ev_before_g1b, st);
        if (e != cudaSuccess) return e;
    }
    const int n = (d >> 4) << 4;      // dims fvec_L2sqr uses
    // the ring of the transposing kernel must fit shared memory: 4 warps x 2 tiles x 16 rows x (n + 8) floats
    if ((d & 3) == 0 && n >= 16 && (size_t)G1B_WARPS * 2 * G1B_TILE * (n + 8) * 4 + (size_t)n * 4 <= 200 * 1024) {
        if (n == 128) e = launch_g1b<128>(st, nq, d, n, queries, L, Wk);
        else if (n == 96) e = launch_g1b<96>(st, nq, d, n, queries, L, Wk);
        else e = launch_g1b<0>(st, nq, d, n, queries, L, Wk);
    } else {
        group_g1b_simple_kernel<<<nq, 256, (size_t)d * 4, st>>>(d, queries, L, Wk);
        e = cudaGetLastError();
    }
    ++*launches;
    if (e != cudaSuccess) return e;
    if (ev_after_g1) {
        e = cudaEventRecord(ev_after_g1, st);
        if (e != cudaSuccess) return e;
    }
    const size_t smc = (size_t)nprobe * 21 + 16;
    group_g1c_kernel<<<nq, G1_THREADS, smc, st>>>(nprobe, nsubc, sub_shift, max_codes, do_pruning ? 1 : 0, coarse_d,
                                                 coarse_i, L, T, Wk);
    ++*launches;
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    if (do_pruning) {
        group_g2_kernel<<<(nq + 31) / 32, 128, 0, st>>>(nq, npairs, T.qsd, T.nscanned, T.threshold);
        ++*launches;
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    const size_t sm3 = (size_t)nprobe * 13 + 16;
    group_g3_kernel<<<nq, G1_THREADS, sm3, st>>>(nprobe, nsubc, sub_shift, max_codes, do_pruning ? 1 : 0, coarse_i, L, T);
    ++*launches;
    return cudaGetLastError();
}

// =============================================================================================
// K-L1: per-query lookup tables, compact form lutc[q][j][m] = -2 <q_m, c_{m,j}>
// (ProductQuantizer::compute_inner_prod_table, faiss/ProductQuantizer.cpp:424-436, scaled by -2 so
//  that the scan adds instead of computing 2*pq_L2sqr, IndexIVF_HNSW.cpp:283; x2 is exact in fp32).
// Thread j owns codeword row j of every sub-quantizer: its reads of the codebook are contiguous across
// the CTA and the M values of a row leave as 128-bit stores.  The codebook (d KB) stays L1 resident
// while a CTA works through its queries.
// =============================================================================================
#define LUT_QPB 8
template <int M>
__global__ void __launch_bounds__(256) lut_kernel(int nq, int d, const float *__restrict__ queries,
                                                  const float *__restrict__ pqc, float *__restrict__ lutc,
                                                  const uint32_t *__restrict__ qnp /* may be null: 0 = skip the query */)
{
    extern __shared__ __align__(16) float sq[];   // [LUT_QPB][d]
    const int q0 = blockIdx.x * LUT_QPB, j = threadIdx.x;
    const int nqb = min(LUT_QPB, nq - q0);
    for (int i = threadIdx.x; i < nqb * d; i += 256) sq[i] = queries[(size_t)q0 * d + i];
    __syncthreads();
    const int dsub = d / M;
    for (int qq = 0; qq < nqb; qq++) {
        if (qnp && qnp[q0 + qq] == 0) continue;     // sharded scan: nothing of this query lives on this rank
        float v[M];
#pragma unroll
        for (int m = 0; m < M; m++) {
            const float *c = pqc + ((size_t)m * 256 + j) * dsub;
            const float *qs = sq + qq * d + m * dsub;
            float s = 0.f;
            if ((dsub & 3) == 0) {
                for (int t = 0; t < dsub; t += 4) {
                    const float4 cv = __ldg(reinterpret_cast<const float4 *>(c + t));
                    s = fmaf(qs[t], cv.x, s);
                    s = fmaf(qs[t + 1], cv.y, s);
                    s = fmaf(qs[t + 2], cv.z, s);
                    s = fmaf(qs[t + 3], cv.w, s);
                }
            } else if ((dsub & 1) == 0) {
                for (int t = 0; t < dsub; t += 2) {
                    const float2 cv = __ldg(reinterpret_cast<const float2 *>(c + t));
                    s = fmaf(qs[t], cv.x, s);
                    s = fmaf(qs[t + 1], cv.y, s);
                }
            } else {
                for (int t = 0; t < dsub; t++) s = fmaf(qs[t], __ldg(c + t), s);
            }
            v[m] = -2.f * s;
        }
        float4 *o = reinterpret_cast<float4 *>(lutc + ((size_t)(q0 + qq) * 256 + j) * M);
#pragma unroll
        for (int m4 = 0; m4 < M / 4; m4++) o[m4] = make_float4(v[m4 * 4], v[m4 * 4 + 1], v[m4 * 4 + 2], v[m4 * 4 + 3]);
    }
}

// Same table, codebook rows held in REGISTERS: thread j loads c_{m,j} for every m once (d floats) and then
// streams LUT_REG_QPB queries through them (query values broadcast from shared memory with 128-bit loads).
// Per query and thread: d FMAs, d/4 LDS.128, M/4 STG.128 -- nothing else.  Same fmaf order as lut_kernel.
#define LUT_REG_QPB 32
template <int M, int DSUB>
__global__ void __launch_bounds__(256) lut_reg_kernel(int nq, const float *__restrict__ queries,
                                                      const float *__restrict__ pqc, float *__restrict__ lutc,
                                                      const uint32_t *__restrict__ qnp /* may be null: 0 = skip */)
{
    constexpr int D = M * DSUB;
    __shared__ __align__(16) float sq[LUT_REG_QPB * D];
    const int q0 = blockIdx.x * LUT_REG_QPB, j = threadIdx.x;
    const int nqb = min(LUT_REG_QPB, nq - q0);
    for (int i = threadIdx.x; i < nqb * D; i += 256) sq[i] = queries[(size_t)q0 * D + i];
    float c[D];
#pragma unroll
    for (int m = 0; m < M; m++)
#pragma unroll
        for (int t = 0; t < DSUB; t++) c[m * DSUB + t] = __ldg(pqc + ((size_t)m * 256 + j) * DSUB + t);
    __syncthreads();
    for (int qq = 0; qq < nqb; qq++) {
        if (qnp && qnp[q0 + qq] == 0) continue;     // sharded scan: nothing of this query lives on this rank
        float qv[D];
#pragma unroll
        for (int i = 0; i < D / 4; i++) {
            const float4 v = *reinterpret_cast<const float4 *>(sq + qq * D + i * 4);
            qv[i * 4] = v.x; qv[i * 4 + 1] = v.y; qv[i * 4 + 2] = v.z; qv[i * 4 + 3] = v.w;
        }
        float4 *o = reinterpret_cast<float4 *>(lutc + ((size_t)(q0 + qq) * 256 + j) * M);
#pragma unroll
        for (int m4 = 0; m4 < M / 4; m4++) {
            float r[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int m = m4 * 4 + u;
                float s = 0.f;
#pragma unroll
                for (int t = 0; t < DSUB; t++) s = fmaf(qv[m * DSUB + t], c[m * DSUB + t], s);
                r[u] = -2.f * s;
            }
            o[m4] = make_float4(r[0], r[1], r[2], r[3]);
        }
    }
}

template <int M, int DSUB>
static cudaError_t launch_lut_reg(cudaStream_t st, int nq, const float *queries, const float *pqc, float *lutc,
                                  const uint32_t *qnp)
{
    lut_reg_kernel<M, DSUB><<<(unsigned)((nq + LUT_REG_QPB - 1) / LUT_REG_QPB), 256, 0, st>>>(nq, queries, pqc, lutc, qnp);
    return cudaGetLastError();
}

cudaError_t launch_lut(cudaStream_t st, uint64_t *launches, int M, int nq, int d, const float *queries,
                       const float *pqc, float *lutc, const uint32_t *qnp)
{
    if (nq == 0) return cudaSuccess;
    ++*launches;
    const int dsub = d / M;
    // register-resident codebook for the usual shapes (d <= 128, d % 4 == 0)
#define IVF_LUT_CASE(MM, DS) if (M == MM && dsub == DS) return launch_lut_reg<MM, DS>(st, nq, queries, pqc, lutc, qnp);
    IVF_LUT_CASE(8, 4) IVF_LUT_CASE(8, 12) IVF_LUT_CASE(8, 16)
    IVF_LUT_CASE(16, 2) IVF_LUT_CASE(16, 6) IVF_LUT_CASE(16, 8)
    IVF_LUT_CASE(32, 3) IVF_LUT_CASE(32, 4)
    IVF_LUT_CASE(64, 2)
#undef IVF_LUT_CASE
    const unsigned grid = (unsigned)((nq + LUT_QPB - 1) / LUT_QPB);
    const size_t sm = (size_t)LUT_QPB * d * 4;
    switch (M) {
    case 8: lut_kernel<8><<<grid, 256, sm, st>>>(nq, d, queries, pqc, lutc, qnp); break;
    case 16: lut_kernel<16><<<grid, 256, sm, st>>>(nq, d, queries, pqc, lutc, qnp); break;
    case 32: lut_kernel<32><<<grid, 256, sm, st>>>(nq, d, queries, pqc, lutc, qnp); break;
    case 64: lut_kernel<64><<<grid, 256, sm, st>>>(nq, d, queries, pqc, lutc, qnp); break;
    default: --*launches; return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

// =============================================================================================
// K-S1/K-S2: the scan
// =============================================================================================
struct ScanArgs {
    int d, nprobe, nsubc, k, nsplit;
    int seg;                  // most list positions per work item (<= SCAN_SEG)
    int seg_min;              // fewest; a query's item length is the power of two in between that gives ~items_target items
    int items_target;
    int pruned;               // Grouping: sub-groups may be pruned (sub_base = +inf) -> tile kept runs only
    const float *lutc;        // [nq][256][M] compact tables from K-L1
    const float *norm_table;  // [256]
    DeviceLists L;
    SegTables T;
    uint32_t c_begin, c_end;  // centroid range whose lists this GPU owns
    float *part_d;            // [nq][nsplit][k]
    uint32_t *part_seq;
    uint32_t *part_slot;
    float *out_d;             // [nq][k]   (nsplit == 1)
    int64_t *out_labels;
    uint32_t *out_seq;        // [nq][k] scan-order keys of the results (optional, sharded search)
};

template <int M> struct ScanCfg {
    static constexpr int ML = M < 32 ? M : 32;      // sub-quantizers resolved by the lane XOR
    static constexpr int R = 32 / ML;               // table copies per sub-quantizer
    static constexpr int NW = M / 4;                // 32-bit words per code
    static constexpr int NO = ML / 4;               // packed per-lane offset registers
    static constexpr int LUT_BYTES = 65536 + (M == 64 ? 32768 : 0);
};

// entries per lane of the register-resident per-warp result list: k <= 32 -> 1, k <= 128 -> 4,
// larger k -> 0 = shared-memory list under a CTA lock
static __host__ __device__ inline int scan_kpl(int k) { return k <= 32 ? 1 : (k <= 128 ? 4 : 0); }

static __host__ __device__ inline size_t scan_small_smem(int nprobe, int k, int nsubc)
{
    size_t b = 0;
    b += (size_t)(6 * nprobe + 1 + SCAN_ITEMS) * 4;   // segment tables, item table
    if (scan_kpl(k)) {
        b += (size_t)SCAN_WARPS * k * 12;     // the warps' sorted lists, for the final merge
        if (scan_kpl(k) == 1) b += (size_t)SCAN_WARPS * 32 * 8;   // warp candidate batches (SCAN_BCAP)
    } else {
        b += (size_t)k * 12;                        // top-k
        b += (size_t)SCAN_WARPS * SCAN_WCAP * 12;   // warp candidate buffers
    }
    if (nsubc > 64) b += (size_t)SCAN_WARPS * nsubc * 4;   // warp sub-group bases (<= 64 live in registers)
    b += (size_t)SCAN_WARPS * SCAN_WBUF * 4;    // warp window lists
    b += 64;                                    // control words
    return (b + 15) & ~(size_t)15;
}

size_t scan_smem_bytes(int M, int nprobe, int k, int nsubc)
{
    return (size_t)(M == 64 ? 98304 : 65536) + scan_small_smem(nprobe, k, nsubc);
}

// control words
struct ScanCtl {
    int piece_ctr;
    int lock;
    int worst_pos;
    uint32_t worst_seq;
    float worst_d;      // == the insertion threshold (lock path)
    uint32_t thr_key;   // min over the warps of fkey(k-th best): a bound on the CTA's final k-th best
    int seg_len;        // list positions per work item of THIS query (power of two, <= SCAN_SEG)
    int pad[1];
};

static __device__ __forceinline__ uint32_t fkey_u32(float f)
{
    const uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);   // order-preserving float -> uint
}

// ---------------------------------------------------------------------------------------------
// Per-warp result list in registers: entry e = r * 32 + lane, ascending (distance, scan order).
// An insertion is one ballot per register row plus a shift by one lane -- no shared memory, no lock.
// ---------------------------------------------------------------------------------------------
template <int KPL> struct WarpList {
    float d[KPL > 0 ? KPL : 1];
    uint32_t s[KPL > 0 ? KPL : 1], slot[KPL > 0 ? KPL : 1];
    float thr_d;      // key of entry k-1: a candidate must come strictly before it
    uint32_t thr_s;

    __device__ __forceinline__ void init(int lane, int k)
    {
#pragma unroll
        for (int r = 0; r < KPL; r++) {
            d[r] = FLT_MAX;
            s[r] = 0xffffff00u + (uint32_t)(r * 32 + lane);   // empty entries: ascending keys above any scan order
            slot[r] = IVF_INVALID_SLOT;
        }
        thr_d = FLT_MAX;
        thr_s = 0xffffff00u + (uint32_t)(k - 1);
    }
    // warp-uniform arguments
    __device__ __forceinline__ void insert(float xd, uint32_t xs, uint32_t xslot, int lane, int k)
    {
        int pos = 0;
#pragma unroll
        for (int r = 0; r < KPL; r++) pos += __popc(__ballot_sync(IVF_FULL_MASK, key_less(d[r], s[r], xd, xs)));
#pragma unroll
        for (int r = KPL - 1; r >= 0; r--) {
            const int e = r * 32 + lane;
            float ud = __shfl_up_sync(IVF_FULL_MASK, d[r], 1);
            uint32_t us = __shfl_up_sync(IVF_FULL_MASK, s[r], 1);
            uint32_t ul = __shfl_up_sync(IVF_FULL_MASK, slot[r], 1);
            if (r > 0) {
                const float cd = __shfl_sync(IVF_FULL_MASK, d[r - 1], 31);
                const uint32_t cs = __shfl_sync(IVF_FULL_MASK, s[r - 1], 31);
                const uint32_t cl = __shfl_sync(IVF_FULL_MASK, slot[r - 1], 31);
                if (lane == 0) { ud = cd; us = cs; ul = cl; }
            }
            if (e == pos) { d[r] = xd; s[r] = xs; slot[r] = xslot; }
            else if (e > pos) { d[r] = ud; s[r] = us; slot[r] = ul; }
        }
        const int kr = (k - 1) >> 5, kl = (k - 1) & 31;
        float td = d[0];
        uint32_t ts = s[0];
#pragma unroll
        for (int r = 1; r < KPL; r++)
            if (kr == r) { td = d[r]; ts = s[r]; }
        thr_d = __shfl_sync(IVF_FULL_MASK, td, kl);
        thr_s = __shfl_sync(IVF_FULL_MASK, ts, kl);
    }
};

// k <= 32: the warp list is one entry per lane and candidates are merged a batch at a time.  Lanes append passing
// candidates to a 32-entry per-warp buffer (one ballot per chunk, nothing serial); a full buffer is sorted by a
// bitonic network in registers, reversed, min-ed against the list and bitonic-merged: ~200 warp instructions
// per 32 candidates instead of ~40 per candidate.  The slot of an entry is recovered from its scan-order key
// at the end, so the network carries (distance, scan order) only.
#define SCAN_BCAP 32
struct BatchList {
    float d, thr_d;
    uint32_t s, thr_s;

    __device__ __forceinline__ void init(int lane, int k)
    {
        d = FLT_MAX;
        s = 0xffffff00u + (uint32_t)lane;
        thr_d = FLT_MAX;
        thr_s = 0xffffff00u + (uint32_t)(k - 1);
    }
    // compare-exchange with lane ^ j: keep the smaller key when keep_small, else the larger
    static __device__ __forceinline__ void cmpx(float &xd, uint32_t &xs, int j, bool keep_small)
    {
        const float od = __shfl_xor_sync(IVF_FULL_MASK, xd, j);
        const uint32_t os = __shfl_xor_sync(IVF_FULL_MASK, xs, j);
        if (key_less(od, os, xd, xs) == keep_small) { xd = od; xs = os; }
    }
    // cd/cs: one candidate per lane, padding = (FLT_MAX, 0xffffffff)
    __device__ __forceinline__ void merge(float cd, uint32_t cs, int lane, int k)
    {
#pragma unroll
        for (int size = 2; size <= 32; size <<= 1) {
#pragma unroll
            for (int j = size >> 1; j > 0; j >>= 1)
                cmpx(cd, cs, j, ((lane & j) == 0) == ((lane & size) == 0));
        }
        // ascending candidates reversed against the ascending list: the lane-wise minimum is the 32 smallest of
        // the 64 keys as a bitonic sequence
        const float rd = __shfl_sync(IVF_FULL_MASK, cd, 31 - lane);
        const uint32_t rs = __shfl_sync(IVF_FULL_MASK, cs, 31 - lane);
        if (key_less(rd, rs, d, s)) { d = rd; s = rs; }
#pragma unroll
        for (int j = 16; j > 0; j >>= 1) cmpx(d, s, j, (lane & j) == 0);
        thr_d = __shfl_sync(IVF_FULL_MASK, d, k - 1);
        thr_s = __shfl_sync(IVF_FULL_MASK, s, k - 1);
    }
    __device__ __forceinline__ void flush(const float *bd, const uint32_t *bs, int count, int lane, int k)
    {
        __syncwarp();
        const bool h = lane < count;
        merge(h ? bd[lane] : FLT_MAX, h ? bs[lane] : 0xffffffffu, lane, k);
        __syncwarp();
    }
};

// Lock path (k > 128): merge one warp's candidate buffer (<= SCAN_WCAP = 64 entries) into the CTA
// top-k.  Warp-collective.  Candidates are pre-filtered lane-parallel against the current worst entry
// (two per lane); only the survivors are inserted one by one under the CTA lock.
static __device__ __forceinline__ void flush_candidates(volatile ScanCtl *ctl, float *topk_d, uint32_t *topk_seq,
                                                        uint32_t *topk_slot, int k, const float *wd,
                                                        const uint32_t *wseq, const uint32_t *wslot, int count,
                                                        int lane)
{
    if (count == 0) return;
    __syncwarp();
    const bool h0 = lane < count, h1 = lane + 32 < count;
    const float d0 = h0 ? wd[lane] : FLT_MAX, d1 = h1 ? wd[lane + 32] : FLT_MAX;
    const uint32_t s0 = h0 ? wseq[lane] : 0xffffffffu, s1 = h1 ? wseq[lane + 32] : 0xffffffffu;
    float w_d = ctl->worst_d;          // unlocked read: stale values are only larger
    uint32_t w_s = ctl->worst_seq;
    bool a0 = h0 && key_less(d0, s0, w_d, w_s), a1 = h1 && key_less(d1, s1, w_d, w_s);
    if (!__any_sync(IVF_FULL_MASK, a0 || a1)) return;
    if (lane == 0) {
        while (atomicCAS((int *)&ctl->lock, 0, 1) != 0) __nanosleep(100);
        __threadfence_block();
    }
    __syncwarp();
    w_d = ctl->worst_d;
    w_s = ctl->worst_seq;
    int wp = ctl->worst_pos;
    while (true) {
        a0 = a0 && key_less(d0, s0, w_d, w_s);
        a1 = a1 && key_less(d1, s1, w_d, w_s);
        const uint32_t b0 = __ballot_sync(IVF_FULL_MASK, a0), b1 = __ballot_sync(IVF_FULL_MASK, a1);
        if (!(b0 | b1)) break;
        const int src = b0 ? (__ffs(b0) - 1) : (__ffs(b1) - 1);
        const float cd = __shfl_sync(IVF_FULL_MASK, b0 ? d0 : d1, src);
        const uint32_t cs = __shfl_sync(IVF_FULL_MASK, b0 ? s0 : s1, src);
        if (lane == src) { if (b0) a0 = false; else a1 = false; }
        if (lane == 0) {
            topk_d[wp] = cd;
            topk_seq[wp] = cs;
            topk_slot[wp] = wslot[b0 ? src : src + 32];
        }
        __syncwarp();
        // new worst = lexicographic max of (dist, seq) over the k entries
        uint32_t bk = 0, bs = 0;
        float bd = 0.f;
        int bp = -1;
        for (int e = lane; e < k; e += 32) {
            const float ed = topk_d[e];
            const uint32_t ek = fkey_u32(ed), es = topk_seq[e];
            if (bp < 0 || ek > bk || (ek == bk && es > bs)) { bk = ek; bs = es; bd = ed; bp = e; }
        }
        const uint32_t mk = __reduce_max_sync(IVF_FULL_MASK, bp >= 0 ? bk : 0u);
        bool is = bp >= 0 && bk == mk;
        const uint32_t ms = __reduce_max_sync(IVF_FULL_MASK, is ? bs : 0u);
        is = is && bs == ms;
        const int win = __ffs(__ballot_sync(IVF_FULL_MASK, is)) - 1;
        wp = __shfl_sync(IVF_FULL_MASK, bp, win);
        w_d = __shfl_sync(IVF_FULL_MASK, bd, win);
        w_s = ms;
    }
    if (lane == 0) {
        ctl->worst_pos = wp;
        ctl->worst_seq = w_s;
        ctl->worst_d = w_d;
        __threadfence_block();
        atomicExch((int *)&ctl->lock, 0);
    }
    __syncwarp();
}

template <int M, bool GROUPING, int KPL>
__global__ void __launch_bounds__(SCAN_THREADS) scan_kernel(ScanArgs a)
{
    using C = ScanCfg<M>;
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t *lut = smem;  // [256][64] floats (+ [256][32] norm replicas for M == 64)
    uint8_t *p = smem + C::LUT_BYTES;
    uint32_t *s_slot = reinterpret_cast<uint32_t *>(p); p += a.nprobe * 4;
    uint32_t *s_n = reinterpret_cast<uint32_t *>(p); p += a.nprobe * 4;
    float *s_base = reinterpret_cast<float *>(p); p += a.nprobe * 4;
    uint32_t *s_seq = reinterpret_cast<uint32_t *>(p); p += a.nprobe * 4;
    uint32_t *s_poff = reinterpret_cast<uint32_t *>(p); p += (a.nprobe + 1) * 4;
    uint32_t *s_cid = reinterpret_cast<uint32_t *>(p); p += a.nprobe * 4;
    uint32_t *s_item = reinterpret_cast<uint32_t *>(p); p += SCAN_ITEMS * 4;
    // KPL > 0: [SCAN_WARPS][k] sorted warp lists; KPL == 0: CTA top-k followed by the warp buffers
    const int nres = KPL > 0 ? SCAN_WARPS * a.k : a.k;
    float *topk_d = reinterpret_cast<float *>(p); p += nres * 4;
    uint32_t *topk_seq = reinterpret_cast<uint32_t *>(p); p += nres * 4;
    uint32_t *topk_slot = reinterpret_cast<uint32_t *>(p); p += nres * 4;
    float *wbuf_d = nullptr;
    uint32_t *wbuf_seq = nullptr, *wbuf_slot = nullptr;
    if (KPL == 0) {
        wbuf_d = reinterpret_cast<float *>(p); p += SCAN_WARPS * SCAN_WCAP * 4;
        wbuf_seq = reinterpret_cast<uint32_t *>(p); p += SCAN_WARPS * SCAN_WCAP * 4;
        wbuf_slot = reinterpret_cast<uint32_t *>(p); p += SCAN_WARPS * SCAN_WCAP * 4;
    } else if (KPL == 1) {
        wbuf_d = reinterpret_cast<float *>(p); p += SCAN_WARPS * SCAN_BCAP * 4;
        wbuf_seq = reinterpret_cast<uint32_t *>(p); p += SCAN_WARPS * SCAN_BCAP * 4;
    }
    float *wsub_all = reinterpret_cast<float *>(p); p += (a.nsubc > 64 ? SCAN_WARPS * a.nsubc * 4 : 0);
    uint32_t *wbuf_all = reinterpret_cast<uint32_t *>(p); p += SCAN_WARPS * SCAN_WBUF * 4;
    p = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(p) + 15) & ~(uintptr_t)15);
    volatile ScanCtl *ctl = reinterpret_cast<volatile ScanCtl *>(p);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int q = blockIdx.x / a.nsplit, split = blockIdx.x - q * a.nsplit;
    const int k = a.k;
    // fixed-stride record of a.nprobe probes, or a packed one holding only the probes this rank owns
    const int nprobe = a.T.qnp ? (int)a.T.qnp[q] : a.nprobe;
    const uint32_t *rec = a.T.qoff ? a.T.rec + a.T.qoff[q] : a.T.rec + (size_t)q * a.T.stride;
    const float *rec_sub_base = reinterpret_cast<const float *>(rec + 5 * (size_t)nprobe);
    if (a.T.qnp && nprobe == 0) {
        // packed records, nothing of this query lives on this rank: the empty result (the order the general path gives
        // k empty entries), before any table is built
        for (int r = tid; r < k; r += SCAN_THREADS) {
            const uint32_t es = 0xffffffffu - (uint32_t)(k - 1 - r);
            if (a.nsplit == 1) {
                a.out_d[(size_t)q * k + r] = FLT_MAX;
                a.out_labels[(size_t)q * k + r] = -1;
                if (a.out_seq) a.out_seq[(size_t)q * k + r] = es;
            } else {
                const size_t o = ((size_t)q * a.nsplit + split) * k + r;
                a.part_d[o] = FLT_MAX;
                a.part_seq[o] = es;
                a.part_slot[o] = IVF_INVALID_SLOT;
            }
        }
        return;
    }

    // ---- stage the query's segment tables (keeping only the lists this GPU owns) and initialise the top-k
    {
        const uint32_t *r_n = rec + nprobe, *r_seq = rec + 3 * (size_t)nprobe, *r_cid = rec + 4 * (size_t)nprobe;
        const float *r_base = reinterpret_cast<const float *>(rec + 2 * (size_t)nprobe);
        for (int i = tid; i < nprobe; i += SCAN_THREADS) {
            const uint32_t c = r_cid[i];
            const bool own = c >= a.c_begin && c < a.c_end;
            // the slot comes from THIS rank's layout, not from the record: the record may have been written by another
            // rank, and ranks built with set_shard lay out only their own lists (slot spaces differ)
            s_slot[i] = own ? a.L.list_slot[c] : 0u;
            s_n[i] = own ? r_n[i] : 0u;
            s_base[i] = r_base[i];
            s_seq[i] = r_seq[i];
            s_cid[i] = c;
        }
        if (KPL == 0) {
            for (int e = tid; e < k; e += SCAN_THREADS) {
                topk_d[e] = FLT_MAX;
                topk_seq[e] = 0xffffffffu - (uint32_t)e;
                topk_slot[e] = IVF_INVALID_SLOT;
            }
        }
        if (tid == 0) {
            ctl->piece_ctr = 0;
            ctl->lock = 0;
            ctl->worst_pos = 0;
            ctl->worst_seq = 0xffffffffu;
            ctl->worst_d = FLT_MAX;
            ctl->thr_key = 0xffffffffu;
        }
    }
    __syncthreads();
    // exclusive prefix of the work items (ceil(n / SCAN_SEG) parts) of the kept lists
    if (warp == 0) {
        // work items of this query: the lists are cut so that the CTA has ~4 items per warp -- a query with little to scan
        // (few probes, a 1/N shard of its lists) would otherwise leave most warps at the final barrier while one
        // walks a whole list (43 % of the warp samples of a 1/8-shard scan were barrier waits)
        uint32_t total = 0;
        for (int b = 0; b < nprobe; b += 32) total += (b + lane < nprobe) ? s_n[b + lane] : 0u;
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) total += __shfl_xor_sync(IVF_FULL_MASK, total, off);
        uint32_t sl = (uint32_t)a.seg_min;
        while (sl < (uint32_t)a.seg && sl * (uint32_t)a.items_target < total) sl <<= 1;
        if (lane == 0) ctl->seg_len = (int)sl;
        uint32_t run = 0;
        for (int b = 0; b < nprobe; b += 32) {
            const int i = b + lane;
            const uint32_t pieces = i < nprobe ? (s_n[i] + sl - 1) / sl : 0u;
            uint32_t incl = pieces;
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const uint32_t t = __shfl_up_sync(IVF_FULL_MASK, incl, off);
                if (lane >= off) incl += t;
            }
            if (i < nprobe) s_poff[i] = run + incl - pieces;
            run += __shfl_sync(IVF_FULL_MASK, incl, 31);
        }
        if (lane == 0) s_poff[nprobe] = run;
        __syncwarp();
        // item -> (list, part) table for the first SCAN_ITEMS items (later ones are found by search)
        for (int i = lane; i < nprobe; i += 32) {
            const uint32_t p0 = s_poff[i], p1 = s_poff[i + 1];
            for (uint32_t t = p0; t < p1 && t < SCAN_ITEMS; t++) s_item[t] = (uint32_t)i | ((t - p0) << 16);
        }
    }
    __syncthreads();
    const int npieces = (int)s_poff[nprobe];
    const uint32_t seg_len = (uint32_t)ctl->seg_len;
    const bool has_work = split < npieces;   // CTA-uniform

    // ---- expand the compact table of K-L1 into the bank-conflict-free layout
    if (has_work) {
        const float *lg = a.lutc + (size_t)q * 256 * M;
        float *lutf = reinterpret_cast<float *>(lut);
        if (M >= 32) {
            // R = 1: row j keeps its M words in order (M == 32: first half of the 64-word row)
#pragma unroll 8
            for (int it = 0; it < M / 4; it++) {
                const int e4 = it * 256 + tid;                 // float4 index in [256 * M / 4)
                const float4 v = __ldg(reinterpret_cast<const float4 *>(lg) + e4);
                const int j = (e4 * 4) / M, m = (e4 * 4) % M;
                *reinterpret_cast<float4 *>(lutf + j * 64 + m) = v;
            }
        } else {
            // 32 consecutive elements = R rows of M values; every value is stored R times side by side
#pragma unroll
            for (int it = 0; it < M; it++) {
                const int e = it * 256 + tid;
                const float v = __ldg(lg + e);
                const int j = e / M, m = e % M;
                if (C::R == 2) *reinterpret_cast<float2 *>(lutf + j * 64 + 2 * m) = make_float2(v, v);
                else *reinterpret_cast<float4 *>(lutf + j * 64 + 4 * m) = make_float4(v, v, v, v);
            }
        }
        // decoded norms (norm_pq->decode, IndexIVF_HNSW.cpp:280): one replica per bank
#pragma unroll
        for (int it = 0; it < 8; it++) {
            const int j = warp * 32 + it * 4 + (lane >> 3);
            const float v = __ldg(a.norm_table + j);
            float *dst = (M == 64) ? (lutf + 16384 + j * 32) : (lutf + j * 64 + 32);
            *reinterpret_cast<float4 *>(dst + (lane & 7) * 4) = make_float4(v, v, v, v);
        }
    }
    __syncthreads();

    // ---- per-lane word offsets: step t reads word R*((t%ML) ^ (lane%ML)) + lane/ML
    uint32_t O[C::NO];
#pragma unroll
    for (int w = 0; w < C::NO; w++) {
        uint32_t v = 0;
#pragma unroll
        for (int b = 0; b < 4; b++) {
            const int t = w * 4 + b;
            const uint32_t word = C::R * ((t ^ lane) & (C::ML - 1)) + lane / C::ML;
            v |= (word * 4u) << (8 * b);
        }
        O[w] = v;
    }
    const uint32_t Onorm = 128u + 4u * lane;  // M < 64: byte0 = offset of word 32 + lane, bytes 1..3 = 0

    WarpList<KPL> wl;
    BatchList bl;
    if (KPL == 1) bl.init(lane, k);
    else if (KPL > 0) wl.init(lane, k);
    float *wd = KPL == 0 ? wbuf_d + warp * SCAN_WCAP : (KPL == 1 ? wbuf_d + warp * SCAN_BCAP : nullptr);
    uint32_t *wseq = KPL == 0 ? wbuf_seq + warp * SCAN_WCAP : (KPL == 1 ? wbuf_seq + warp * SCAN_BCAP : nullptr);
    uint32_t *wslot = KPL == 0 ? wbuf_slot + warp * SCAN_WCAP : nullptr;
    float *wsub = wsub_all + warp * a.nsubc;
    int wcount = 0;

    // ---- K-S: pull work items = (list, SCAN_SEG-position part of it).  A part is processed as a sequence of
    // WINDOWS of W consecutive list positions; a window may start at any position: lane l takes the position p
    // of the window with p % W == l % W, so that (slot % ML) == (l % ML) and the table gathers stay conflict
    // free (list starts are aligned).  M <= 16: two 16-wide windows per warp step (one per half warp);
    // M >= 32: one 32-wide window.  Without pruning the windows tile the part; with pruning (Grouping) only
    // the maximal runs of kept sub-groups are tiled, each from its own start, so pruned codes are neither
    // loaded nor evaluated.
    constexpr int W = M >= 32 ? 32 : 16;
    constexpr int WPS = 32 / W;
    const int half = (W == 16) ? (lane >> 4) : 0;
    uint32_t *wbuf = wbuf_all + warp * SCAN_WBUF;
    float thr_f = FLT_MAX;   // prefilter: min(own k-th best, bound shared by the CTA); ties pass, exact test follows
    while (has_work) {
        int pc = 0;
        if (lane == 0) pc = atomicAdd((int *)&ctl->piece_ctr, 1);
        pc = __shfl_sync(IVF_FULL_MASK, pc, 0);
        const int piece = pc * a.nsplit + split;
        if (piece >= npieces) break;
        int seg = 0;
        uint32_t lo;
        if (piece < SCAN_ITEMS) {
            const uint32_t it = s_item[piece];
            seg = (int)(it & 0xffffu);
            lo = (it >> 16) * seg_len;
        } else {
            for (int b = 0; b < nprobe; b += 32) {
                const int i = b + lane;
                const bool le = (i < nprobe) && ((int)s_poff[i] <= piece);
                seg += __popc(__ballot_sync(IVF_FULL_MASK, le));
            }
            seg -= 1;
            lo = (uint32_t)(piece - (int)s_poff[seg]) * seg_len;
        }
        const uint32_t hi = min(s_n[seg], lo + seg_len);
        const uint32_t lslot = s_slot[seg], lseq = s_seq[seg];
        const float base = s_base[seg];
        int nwin = (int)((hi - lo + W - 1) / W);
        bool runs = false;
        // sub-group bases of this list: up to 64 sub-groups live in two registers per lane and are fetched
        // with a shuffle; more go through shared memory
        float sb0 = 0.f, sb1 = 0.f;
        const bool sb_regs = a.nsubc <= 64;
        if (GROUPING) {
            const float pinf = __int_as_float(0x7f800000);
            const float *sbg = rec_sub_base + (size_t)seg * a.nsubc;
            if (sb_regs) {
                sb0 = lane < a.nsubc ? __ldg(sbg + lane) : pinf;
                sb1 = lane + 32 < a.nsubc ? __ldg(sbg + lane + 32) : pinf;
                if (W == 16 && a.pruned) {
                    // positions of the sub-groups inside the list (lane: sub-groups lane and lane + 32)
                    const uint32_t *szg = a.L.sub_sizes + (size_t)s_cid[seg] * a.nsubc;
                    const uint32_t sz0 = lane < a.nsubc ? __ldg(szg + lane) : 0u;
                    const uint32_t sz1 = lane + 32 < a.nsubc ? __ldg(szg + lane + 32) : 0u;
                    uint32_t in0 = sz0, in1 = sz1;
#pragma unroll
                    for (int off = 1; off < 32; off <<= 1) {
                        const uint32_t t0 = __shfl_up_sync(IVF_FULL_MASK, in0, off);
                        const uint32_t t1 = __shfl_up_sync(IVF_FULL_MASK, in1, off);
                        if (lane >= off) { in0 += t0; in1 += t1; }
                    }
                    in1 += __shfl_sync(IVF_FULL_MASK, in0, 31);
                    const uint32_t o0 = max(in0 - sz0, lo), e0 = min(in0, hi);
                    const uint32_t o1 = max(in1 - sz1, lo), e1 = min(in1, hi);
                    const bool ne0 = o0 < e0, ne1 = o1 < e1;             // has codes inside the part
                    const bool k0 = ne0 && sb0 < pinf, k1 = ne1 && sb1 < pinf;   // ... and is kept
                    // run start = end of the last pruned code before the sub-group, run end = first pruned code
                    // after it: nearest set bits of the pruned-and-non-empty mask, bounds fetched by shuffle
                    const uint32_t P_lo = __ballot_sync(IVF_FULL_MASK, ne0 && !k0);
                    const uint32_t P_hi = __ballot_sync(IVF_FULL_MASK, ne1 && !k1);
                    const uint32_t below = (1u << lane) - 1u, above = 0xfffffffeu << lane;
                    const uint32_t bl0 = P_lo & below, bl1 = P_hi & below, ab0 = P_lo & above, ab1 = P_hi & above;
                    const uint32_t rsa = __shfl_sync(IVF_FULL_MASK, e0, 31 - __clz(bl0 | 1u));
                    const uint32_t rsb = __shfl_sync(IVF_FULL_MASK, e1, 31 - __clz(bl1 | 1u));
                    const uint32_t rsc = __shfl_sync(IVF_FULL_MASK, e0, 31 - __clz(P_lo | 1u));
                    const uint32_t rea = __shfl_sync(IVF_FULL_MASK, o0, (__ffs(ab0) - 1) & 31);
                    const uint32_t reb = __shfl_sync(IVF_FULL_MASK, o1, (__ffs(ab1) - 1) & 31);
                    const uint32_t rec = __shfl_sync(IVF_FULL_MASK, o1, (__ffs(P_hi) - 1) & 31);
                    const uint32_t rs0 = bl0 ? rsa : lo;
                    const uint32_t rs1 = bl1 ? rsb : (P_lo ? rsc : lo);
                    const uint32_t re0 = ab0 ? rea : (P_hi ? rec : hi);
                    const uint32_t re1 = ab1 ? reb : hi;
                    // windows of the run's 16-grid that start inside the sub-group
                    const uint32_t f0 = rs0 + ((o0 - rs0 + 15u) & ~15u), f1 = rs1 + ((o1 - rs1 + 15u) & ~15u);
                    const uint32_t c0 = (k0 && f0 < e0) ? ((e0 - f0 + 15u) >> 4) : 0u;
                    const uint32_t c1 = (k1 && f1 < e1) ? ((e1 - f1 + 15u) >> 4) : 0u;
                    uint32_t ic = c0 | (c1 << 16);       // both counts in one scan (<= SCAN_SEG / 16 + 32 each)
#pragma unroll
                    for (int off = 1; off < 32; off <<= 1) {
                        const uint32_t t = __shfl_up_sync(IVF_FULL_MASK, ic, off);
                        if (lane >= off) ic += t;
                    }
                    const uint32_t tot = __shfl_sync(IVF_FULL_MASK, ic, 31);
                    const int total = (int)((tot & 0xffffu) + (tot >> 16));
                    if (total < nwin) {      // else: walking the whole part costs no more
                        runs = true;
                        nwin = total;
                        const uint32_t i0 = (ic & 0xffffu) - c0, i1 = (tot & 0xffffu) + (ic >> 16) - c1;
                        __syncwarp();
                        for (uint32_t t = 0; t < c0; t++) {
                            const uint32_t p0 = f0 + 16u * t;
                            wbuf[i0 + t] = (p0 - lo) | (min(16u, re0 - p0) << 16);
                        }
                        for (uint32_t t = 0; t < c1; t++) {
                            const uint32_t p1 = f1 + 16u * t;
                            wbuf[i1 + t] = (p1 - lo) | (min(16u, re1 - p1) << 16);
                        }
                        __syncwarp();
                    }
                }
            } else {
                __syncwarp();
                for (int s = lane; s < a.nsubc; s += 32) wsub[s] = sbg[s];
                __syncwarp();
            }
        }
        if (!runs) {   // plain tiling of the part: same window list format ((start - lo) | len << 16)
            __syncwarp();
            for (int wi = lane; wi < nwin; wi += 32) {
                const uint32_t rel = (uint32_t)wi * W;
                wbuf[wi] = rel | (min((uint32_t)W, hi - lo - rel) << 16);
            }
            __syncwarp();
        }
        if (KPL == 1 || KPL > 0) {
            const uint32_t sh = ctl->thr_key;
            const uint32_t fb = (sh & 0x80000000u) ? (sh ^ 0x80000000u) : ~sh;   // inverse of fkey_u32
            thr_f = fminf(KPL == 1 ? bl.thr_d : wl.thr_d, __uint_as_float(fb));
        }

        // a batch of NS warp steps: every load is issued before the first use
        auto batch = [&](auto ns_tag, const int j0, const int nlive) {
            constexpr int NS = decltype(ns_tag)::value;
            uint32_t cw[NS][C::NW];
            uint32_t nm[NS], sid[NS], posv[NS];
#pragma unroll
            for (int c = 0; c < NS; c++) {
                const int wi = (j0 + c) * WPS + half;
                uint32_t start = 0, len = 0;
                if (c < nlive && wi < nwin) {
                    const uint32_t w = wbuf[wi];
                    start = lo + (w & 0xffffu);
                    len = w >> 16;
                }
                const uint32_t rot = ((uint32_t)lane - start) & (W - 1);
                const bool valid = rot < len;
                posv[c] = valid ? start + rot : 0xffffffffu;
                nm[c] = 0;
                sid[c] = 0;
                if (valid) {
                    const uint32_t slot = lslot + start + rot;
                    const uint8_t *cp = a.L.codes + (size_t)slot * M;
                    if (M == 8) {
                        const uint2 v = ldg_stream_64(cp);
                        cw[c][0] = v.x; cw[c][1] = v.y;
                    } else {
#pragma unroll
                        for (int w4 = 0; w4 < C::NW / 4; w4++) {
                            const uint4 v = ldg_stream_128(cp + w4 * 16);
                            cw[c][w4 * 4 + 0] = v.x; cw[c][w4 * 4 + 1] = v.y;
                            cw[c][w4 * 4 + 2] = v.z; cw[c][w4 * 4 + 3] = v.w;
                        }
                    }
                    nm[c] = ldg_stream_u8(a.L.norm_codes + slot);
                    if (GROUPING) sid[c] = ldg_stream_u8(a.L.sub_ids + slot);
                }
            }
#pragma unroll
            for (int c = 0; c < NS; c++) {
                if (c < nlive) {   // warp-uniform
                    float acc0 = base;
                    if (GROUPING) {
                        if (sb_regs) {
                            const float v0 = __shfl_sync(IVF_FULL_MASK, sb0, sid[c] & 31);
                            const float v1 = __shfl_sync(IVF_FULL_MASK, sb1, sid[c] & 31);
                            acc0 = (sid[c] & 32) ? v1 : v0;
                        } else acc0 = wsub[sid[c]];
                    }
                    float acc1;
                    if (M == 64) acc1 = *reinterpret_cast<const float *>(lut + 65536 + nm[c] * 128u + 4u * lane);
                    else acc1 = *reinterpret_cast<const float *>(lut + prmt(nm[c], Onorm, 0x5504u));
#pragma unroll
                    for (int t = 0; t < M; t++) {
                        const int b = t & 3;
                        const uint32_t sel = 0xCC00u | (uint32_t)(b << 4) | (uint32_t)(4 + b);
                        const uint32_t addr = prmt(cw[c][t >> 2], O[(t & (C::ML - 1)) >> 2], sel);
                        const float v = *reinterpret_cast<const float *>(lut + (t >= 32 ? 128 : 0) + addr);
                        if (t & 1) acc1 += v; else acc0 += v;
                    }
                    const float dist = acc0 + acc1;
                    const bool valid = posv[c] != 0xffffffffu;
                    if (KPL == 1) {
                        if (__any_sync(IVF_FULL_MASK, valid && dist <= thr_f)) {
                            const uint32_t myseq = lseq + posv[c];
                            bool pass = valid && dist <= thr_f && key_less(dist, myseq, bl.thr_d, bl.thr_s);
                            uint32_t bal = __ballot_sync(IVF_FULL_MASK, pass);
                            if (bal) {
                                int n = __popc(bal);
                                if (wcount + n > SCAN_BCAP) {
                                    bl.flush(wd, wseq, wcount, lane, k);
                                    wcount = 0;
                                    if (lane == 0 && bl.thr_d < FLT_MAX)
                                        atomicMin((uint32_t *)&ctl->thr_key, fkey_u32(bl.thr_d));
                                    thr_f = fminf(thr_f, bl.thr_d);
                                    pass = pass && key_less(dist, myseq, bl.thr_d, bl.thr_s);
                                    bal = __ballot_sync(IVF_FULL_MASK, pass);
                                    n = __popc(bal);
                                }
                                if (pass) {
                                    const int pos = wcount + __popc(bal & ((1u << lane) - 1u));
                                    wd[pos] = dist;
                                    wseq[pos] = myseq;
                                }
                                wcount += n;
                            }
                        }
                    } else if (KPL > 0) {
                        uint32_t bal = __ballot_sync(IVF_FULL_MASK, valid && dist <= thr_f);
                        if (bal) {
                            do {
                                const int src = __ffs(bal) - 1;
                                bal &= bal - 1;
                                const float xd = __shfl_sync(IVF_FULL_MASK, dist, src);
                                const uint32_t xp = __shfl_sync(IVF_FULL_MASK, posv[c], src);
                                if (key_less(xd, lseq + xp, wl.thr_d, wl.thr_s))   // warp-uniform
                                    wl.insert(xd, lseq + xp, lslot + xp, lane, k);
                            } while (bal);
                            if (lane == 0 && wl.thr_d < FLT_MAX) atomicMin((uint32_t *)&ctl->thr_key, fkey_u32(wl.thr_d));
                            thr_f = fminf(thr_f, wl.thr_d);
                        }
                    } else {
                        const float thr = ctl->worst_d;
                        bool pass = valid && dist <= thr;
                        uint32_t bal = __ballot_sync(IVF_FULL_MASK, pass);
                        if (bal) {
                            int n = __popc(bal);
                            if (wcount + n > SCAN_WCAP) {
                                flush_candidates(ctl, topk_d, topk_seq, topk_slot, k, wd, wseq, wslot, wcount, lane);
                                wcount = 0;
                                pass = valid && dist <= ctl->worst_d;
                                bal = __ballot_sync(IVF_FULL_MASK, pass);
                                n = __popc(bal);
                            }
                            if (pass) {
                                const int pos = wcount + __popc(bal & ((1u << lane) - 1u));
                                wd[pos] = dist;
                                wseq[pos] = lseq + posv[c];
                                wslot[pos] = lslot + posv[c];
                            }
                            wcount += n;
                        }
                    }
                }
            }
        };
        const int nsteps = (nwin + WPS - 1) / WPS;
        // every batch costs a memory round trip: the tail is one batch whose steps past the end are skipped
        for (int j0 = 0; j0 < nsteps; j0 += SCAN_PF) batch(std::integral_constant<int, SCAN_PF>(), j0, nsteps - j0);
    }

    if (KPL > 0) {
        // ---- publish the warp lists, then rank every entry among the other warps' (sorted) lists
        if (KPL == 1) {
            if (wcount) bl.flush(wd, wseq, wcount, lane, k);
            // empty entries: one ascending key per position, above any scan order
            const bool empty = bl.s >= 0xffffff00u;
            if (lane < k) {
                topk_d[warp * k + lane] = empty ? FLT_MAX : bl.d;
                topk_seq[warp * k + lane] = empty ? 0xffffff00u + (uint32_t)lane : bl.s;
            }
        }
#pragma unroll
        for (int r = 0; r < (KPL == 1 ? 0 : KPL); r++) {
            const int e = r * 32 + lane;
            if (e < k) {
                topk_d[warp * k + e] = wl.d[r];
                topk_seq[warp * k + e] = wl.s[r];
                topk_slot[warp * k + e] = wl.slot[r];
            }
        }
        __syncthreads();
        for (int e = tid; e < SCAN_WARPS * k; e += SCAN_THREADS) {
            const int mine = e / k, pos = e - mine * k;
            const float ed = topk_d[e];
            const uint32_t es = topk_seq[e];
            int rank = pos;
            for (int s = 0; s < SCAN_WARPS && rank < k; s++) {
                if (s == mine) continue;
                const float *ld = topk_d + s * k;
                const uint32_t *ls = topk_seq + s * k;
                int lo = 0, hi = k;
                while (lo < hi) {
                    const int mid = (lo + hi) >> 1;
                    const float md = ld[mid];
                    const uint32_t ms = ls[mid];
                    // empty entries of different warps share keys: break those ties by warp
                    const bool before = key_less(md, ms, ed, es) || (md == ed && ms == es && s < mine);
                    if (before) lo = mid + 1; else hi = mid;
                }
                rank += lo;
            }
            if (rank < k) {
                uint32_t slot;
                if (KPL == 1) {
                    // the list of scan-order key es: the last one starting at or before it
                    slot = IVF_INVALID_SLOT;
                    if (es < 0xffffff00u) {
                        int lo = 0, hi = nprobe;
                        while (lo < hi) {
                            const int mid = (lo + hi) >> 1;
                            if (s_seq[mid] <= es) lo = mid + 1; else hi = mid;
                        }
                        slot = s_slot[lo - 1] + (es - s_seq[lo - 1]);
                    }
                } else slot = topk_slot[e];
                if (a.nsplit == 1) {
                    a.out_d[(size_t)q * k + rank] = ed;
                    a.out_labels[(size_t)q * k + rank] = slot == IVF_INVALID_SLOT ? (int64_t)-1 : (int64_t)a.L.ids[slot];
                    if (a.out_seq) a.out_seq[(size_t)q * k + rank] = es;
                } else {
                    const size_t o = ((size_t)q * a.nsplit + split) * k + rank;
                    a.part_d[o] = ed;
                    a.part_seq[o] = es;
                    a.part_slot[o] = slot;
                }
            }
        }
        return;
    }

    flush_candidates(ctl, topk_d, topk_seq, topk_slot, k, wd, wseq, wslot, wcount, lane);
    __syncthreads();

    // ---- rank-sort the k survivors by (distance, scan order) and write them out
    for (int e = tid; e < k; e += SCAN_THREADS) {
        const float ed = topk_d[e];
        const uint32_t es = topk_seq[e];
        int rank = 0;
        for (int o = 0; o < k; o++) rank += key_less(topk_d[o], topk_seq[o], ed, es) ? 1 : 0;
        const uint32_t slot = topk_slot[e];
        if (a.nsplit == 1) {
            a.out_d[(size_t)q * k + rank] = ed;
            a.out_labels[(size_t)q * k + rank] = slot == IVF_INVALID_SLOT ? (int64_t)-1 : (int64_t)a.L.ids[slot];
            if (a.out_seq) a.out_seq[(size_t)q * k + rank] = es;
        } else {
            const size_t o = ((size_t)q * a.nsplit + split) * k + rank;
            a.part_d[o] = ed;
            a.part_seq[o] = es;
            a.part_slot[o] = slot;
        }
    }
}

template <int M, bool G, int KPL>
static cudaError_t launch_scan_t(cudaStream_t st, int nq, const ScanArgs &a)
{
    const size_t smem = scan_smem_bytes(M, a.nprobe, a.k, a.nsubc);
    cudaError_t e =
        cudaFuncSetAttribute(scan_kernel<M, G, KPL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    scan_kernel<M, G, KPL><<<(unsigned)((size_t)nq * a.nsplit), SCAN_THREADS, smem, st>>>(a);
    return cudaGetLastError();
}

template <int M, bool G>
static cudaError_t launch_scan_k(cudaStream_t st, int nq, const ScanArgs &a)
{
    switch (scan_kpl(a.k)) {
    case 1: return launch_scan_t<M, G, 1>(st, nq, a);
    case 4: return launch_scan_t<M, G, 4>(st, nq, a);
    default: return launch_scan_t<M, G, 0>(st, nq, a);
    }
}

cudaError_t launch_scan(cudaStream_t st, uint64_t *launches, int M, bool grouping, int nq, int d, int nprobe,
                        int nsubc, bool pruned, int k, int nsplit, const float *lutc, const float *norm_table,
                        const DeviceLists &L, const SegTables &T, uint32_t c_begin, uint32_t c_end, float *part_d,
                        uint32_t *part_seq, uint32_t *part_slot, float *out_d, int64_t *out_labels, uint32_t *out_seq)
{
    if (nq == 0) return cudaSuccess;
    ScanArgs a;
    a.out_seq = out_seq;
    a.pruned = (grouping && pruned) ? 1 : 0;
    // few probes: cut the lists finer so that all 8 warps of a query's CTA find work (ENN: ONE list per query)
    a.seg = nprobe >= 8 ? SCAN_SEG : 128;
    a.seg_min = 128;
    static const int items_env = getenv("IVFHNSW_B200_SCAN_ITEMS") ? atoi(getenv("IVFHNSW_B200_SCAN_ITEMS")) : 0;   // profiling switch
    a.items_target = items_env > 0 ? items_env : 32;
    a.d = d; a.nprobe = nprobe; a.nsubc = grouping ? nsubc : 0; a.k = k; a.nsplit = nsplit;
    a.lutc = lutc; a.norm_table = norm_table; a.L = L; a.T = T; a.c_begin = c_begin; a.c_end = c_end;
    a.part_d = part_d; a.part_seq = part_seq; a.part_slot = part_slot; a.out_d = out_d; a.out_labels = out_labels;
    ++*launches;
#define IVF_SCAN_CASE(MM)                                                              \
    case MM: return grouping ? launch_scan_k<MM, true>(st, nq, a) : launch_scan_k<MM, false>(st, nq, a);
    switch (M) {
        IVF_SCAN_CASE(8)
        IVF_SCAN_CASE(16)
        IVF_SCAN_CASE(32)
        IVF_SCAN_CASE(64)
    }
#undef IVF_SCAN_CASE
    --*launches;
    return cudaErrorInvalidValue;
}

// =============================================================================================
// Sharded search: records -> one packed stream per owning rank (kernels.h launch_pack_records).
// The preparation (coarse stage, pruning pass, max_codes cut) runs on the query's HOME rank; a scanning rank needs,
// per query, only the probes it owns.  Shipping exactly those (an all-to-all-v of ~(5 + nsubc) words per owned probe)
// instead of all-gathering every 35 KB record is what lets the list-sharded layout scale.
// =============================================================================================
#define PACK_MAX_PARTS 64
__global__ void __launch_bounds__(128) pack_count_kernel(int nprobe, int nc_per, int nparts, const uint32_t *__restrict__ rec,
                                                         size_t stride, int nq, uint32_t *__restrict__ cnt)
{
    __shared__ uint32_t s_cnt[PACK_MAX_PARTS];
    const int q = blockIdx.x;
    for (int r = threadIdx.x; r < nparts; r += 128) s_cnt[r] = 0;
    __syncthreads();
    const uint32_t *rn = rec + (size_t)q * stride + nprobe, *rc = rec + (size_t)q * stride + 4 * (size_t)nprobe;
    for (int i = threadIdx.x; i < nprobe; i += 128)
        if (rn[i]) atomicAdd(&s_cnt[min((int)(rc[i] / (uint32_t)nc_per), nparts - 1)], 1u);
    __syncthreads();
    for (int r = threadIdx.x; r < nparts; r += 128) cnt[(size_t)r * nq + q] = s_cnt[r];
}

// one CTA per part: off[r][q] = exclusive scan over q of cnt[r][q] * W; pwords[r] = total
__global__ void __launch_bounds__(1024) pack_scan_kernel(int nq, int W, const uint32_t *__restrict__ cnt,
                                                         uint32_t *__restrict__ off, unsigned long long *__restrict__ pwords)
{
    __shared__ unsigned long long s_v[1024];
    __shared__ unsigned long long carry;
    const int r = blockIdx.x;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < nq; base += 1024) {
        const int q = base + threadIdx.x;
        const unsigned long long w = q < nq ? (unsigned long long)cnt[(size_t)r * nq + q] * W : 0ull;
        s_v[threadIdx.x] = w;
        __syncthreads();
        for (int o = 1; o < 1024; o <<= 1) {
            const unsigned long long t = (int)threadIdx.x >= o ? s_v[threadIdx.x - o] : 0ull;
            __syncthreads();
            s_v[threadIdx.x] += t;
            __syncthreads();
        }
        if (q < nq) off[(size_t)r * nq + q] = (uint32_t)(carry + s_v[threadIdx.x] - w);
        __syncthreads();
        if (threadIdx.x == 1023) carry += s_v[1023];
        __syncthreads();
    }
    if (threadIdx.x == 0) pwords[r] = carry;
}

__global__ void __launch_bounds__(128) pack_fill_kernel(int nprobe, int nsubc, int nc_per, int nparts,
                                                        const uint32_t *__restrict__ rec, size_t stride, int nq,
                                                        const uint32_t *__restrict__ cnt, const uint32_t *__restrict__ off,
                                                        const unsigned long long *__restrict__ pwords,
                                                        uint32_t *__restrict__ out)
{
    extern __shared__ uint32_t pk_sm[];
    uint32_t *s_dst = pk_sm;                 // [nprobe] word offset of the probe's mini record start in `out`, ~0 = dropped
    uint32_t *s_pos = s_dst + nprobe;        // [nprobe] position inside its (query, part)
    uint32_t *s_np = s_pos + nprobe;         // [nprobe] probes of its (query, part)
    __shared__ unsigned long long s_pbase[PACK_MAX_PARTS];
    const int q = blockIdx.x, tid = threadIdx.x, lane = tid & 31;
    const uint32_t *r0 = rec + (size_t)q * stride;
    if (tid == 0) {
        unsigned long long run = 0;
        for (int r = 0; r < nparts; r++) { s_pbase[r] = run; run += pwords[r]; }
    }
    __syncthreads();
    if (tid < 32) {
        // order-preserving position of every kept probe among the probes of its part: ballots per part
        uint32_t run[1];
        for (int r = 0; r < nparts; r++) {
            run[0] = 0;
            const uint32_t np = cnt[(size_t)r * nq + q];
            const unsigned long long start = s_pbase[r] + off[(size_t)r * nq + q];
            for (int b = 0; b < nprobe; b += 32) {
                const int i = b + lane;
                const bool mine = i < nprobe && r0[nprobe + i] != 0u &&
                                  min((int)(r0[4 * (size_t)nprobe + i] / (uint32_t)nc_per), nparts - 1) == r;
                const uint32_t bal = __ballot_sync(IVF_FULL_MASK, mine);
                if (mine) {
                    s_pos[i] = run[0] + __popc(bal & ((1u << lane) - 1u));
                    s_np[i] = np;
                    s_dst[i] = (uint32_t)start;     // part streams are < 2^32 words (checked on the host)
                }
                run[0] += __popc(bal);
            }
        }
        for (int i = lane; i < nprobe; i += 32)
            if (r0[nprobe + i] == 0u) s_dst[i] = 0xffffffffu;
    }
    __syncthreads();
    // the five scalar fields
    for (int e = tid; e < 5 * nprobe; e += 128) {
        const int f = e / nprobe, i = e - f * nprobe;
        if (s_dst[i] != 0xffffffffu) out[(size_t)s_dst[i] + (size_t)f * s_np[i] + s_pos[i]] = r0[(size_t)f * nprobe + i];
    }
    // sub-group bases: coalesced rows of nsubc words
    for (int e = tid; e < nprobe * nsubc; e += 128) {
        const int i = e / nsubc, sg = e - i * nsubc;
        if (s_dst[i] != 0xffffffffu)
            out[(size_t)s_dst[i] + 5 * (size_t)s_np[i] + (size_t)s_pos[i] * nsubc + sg] = r0[5 * (size_t)nprobe + e];
    }
}

cudaError_t launch_pack_records(cudaStream_t st, uint64_t *launches, int nq, int nprobe, int nsubc, int nc, int nparts,
                                const uint32_t *rec, size_t stride, uint32_t *out, uint32_t *cnt, uint32_t *off,
                                unsigned long long *pwords)
{
    if (nq == 0) return cudaSuccess;
    if (nparts < 1 || nparts > PACK_MAX_PARTS) return cudaErrorInvalidValue;
    const int nc_per = (nc + nparts - 1) / nparts;
    pack_count_kernel<<<nq, 128, 0, st>>>(nprobe, nc_per, nparts, rec, stride, nq, cnt);
    pack_scan_kernel<<<nparts, 1024, 0, st>>>(nq, 5 + nsubc, cnt, off, pwords);
    pack_fill_kernel<<<nq, 128, (size_t)nprobe * 12, st>>>(nprobe, nsubc, nc_per, nparts, rec, stride, nq, cnt, off, pwords,
                                                          out);
    *launches += 3;
    return cudaGetLastError();
}

// =============================================================================================
// K-M1: merge the sorted partial top-k lists of each query (keys (dist, seq) are unique per query).
// List s of query q starts at q * q_stride + s * p_stride.  LABELS_IN = false: payload is a slot that
// is translated through ids[] (splits of one GPU); true: payload already is the int64 label (shards
// gathered from several GPUs).
// =============================================================================================
template <bool LABELS_IN>
__global__ void __launch_bounds__(256) merge_kernel(int k, int nparts, size_t q_stride, size_t p_stride,
                                                    const float *__restrict__ part_d,
                                                    const uint32_t *__restrict__ part_seq,
                                                    const void *__restrict__ part_payload,
                                                    const uint32_t *__restrict__ ids, float *out_d,
                                                    int64_t *out_labels, uint32_t *out_seq)
{
    const int q = blockIdx.x;
    const size_t base = (size_t)q * q_stride;
    const int n = nparts * k;
    for (int e = threadIdx.x; e < n; e += blockDim.x) {
        const int mine = e / k, pos = e - mine * k;
        const size_t me = base + (size_t)mine * p_stride + pos;
        const float ed = part_d[me];
        const uint32_t es = part_seq[me];
        int rank = pos;  // elements before it in its own (sorted) list
        for (int s = 0; s < nparts && rank < k; s++) {
            if (s == mine) continue;
            const float *ld = part_d + base + (size_t)s * p_stride;
            const uint32_t *ls = part_seq + base + (size_t)s * p_stride;
            int lo = 0, hi = k;  // count of elements of list s that come before (ed, es)
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                // empty entries of different lists share keys (FLT_MAX, 0xffffffff - e): break ties by list
                const float md = ld[mid];
                const uint32_t ms = ls[mid];
                const bool before = key_less(md, ms, ed, es) || (md == ed && ms == es && s < mine);
                if (before) lo = mid + 1; else hi = mid;
            }
            rank += lo;
        }
        if (rank < k) {
            int64_t lab;
            if (LABELS_IN) lab = reinterpret_cast<const int64_t *>(part_payload)[me];
            else {
                const uint32_t slot = reinterpret_cast<const uint32_t *>(part_payload)[me];
                lab = slot == IVF_INVALID_SLOT ? (int64_t)-1 : (int64_t)ids[slot];
            }
            out_d[(size_t)q * k + rank] = ed;
            out_labels[(size_t)q * k + rank] = lab;
            if (out_seq) out_seq[(size_t)q * k + rank] = es;
        }
    }
}

cudaError_t launch_merge(cudaStream_t st, uint64_t *launches, int nq, int k, int nsplit, const float *part_d,
                         const uint32_t *part_seq, const uint32_t *part_slot, const uint32_t *ids,
                         float *out_d, int64_t *out_labels, uint32_t *out_seq)
{
    if (nq == 0) return cudaSuccess;
    merge_kernel<false><<<nq, 256, 0, st>>>(k, nsplit, (size_t)nsplit * k, (size_t)k, part_d, part_seq, part_slot, ids,
                                            out_d, out_labels, out_seq);
    ++*launches;
    return cudaGetLastError();
}

cudaError_t launch_merge_shards(cudaStream_t st, uint64_t *launches, int nq, int k, int nparts, const float *part_d,
                                const uint32_t *part_seq, const int64_t *part_labels, float *out_d,
                                int64_t *out_labels)
{
    if (nq == 0) return cudaSuccess;
    merge_kernel<true><<<nq, 256, 0, st>>>(k, nparts, (size_t)k, (size_t)nq * k, part_d, part_seq, part_labels, nullptr,
                                           out_d, out_labels, nullptr);
    ++*launches;
    return cudaGetLastError();
}

}  // namespace ivfb200
