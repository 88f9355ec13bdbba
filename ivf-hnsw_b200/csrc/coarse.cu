// Coarse stage (K-A1 / K-A2 / K-G3): exact nearest centroids as a GEMM with a fused
// "best COARSE_T of every group of COARSE_GROUP centroids" epilogue, followed by a per-query
// select + verify + exact fp32 re-rank.
//
// Replaces hnswlib::HierarchicalNSW::searchKnn as used by
//   IndexIVF_HNSW::assign                (IndexIVF_HNSW.cpp:68-72)     nprobe = 1
//   IndexIVF_HNSW::search                (IndexIVF_HNSW.cpp:248-259)   nprobe nearest, ascending
//   IndexIVF_HNSW_Grouping::add_group    (IndexIVF_HNSW_Grouping.cpp:47-63) nsubc nearest other centroids
// with an EXACT search.  Output distances use the reference's fstdistfunc numerics
// (hnswlib/hnswalg.cpp:326-356: floor(d/16)*16 dims, 8 partial sums).
//
// Exactness argument: the GEMM keeps, per (query, group), the COARSE_T smallest approximate
// distances.  Let tau be the K'-th smallest kept value (K' = nprobe + slack).  Any element a group
// did NOT keep is >= that group's COARSE_T-th kept value; so if that value is > tau the group hides
// nothing that could be among the K' best.  Groups that fail the test are expanded in full.
// The K' (+ expansions) candidates are then re-ranked with exact fp32 distances.
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"

namespace ivfb200 {

// ---------------------------------------------------------------------------------------------
// impl 0: fp32 SIMT tile GEMM, 128 queries x 128 centroids per CTA, 8x8 micro-tiles.
// value(q, c) = ||c||^2 - 2 <q, c>   (||q||^2 is constant per query and irrelevant for ranking)
// ---------------------------------------------------------------------------------------------
#define CG_BM 128
#define CG_BN COARSE_GROUP
#define CG_BK 8

__global__ void __launch_bounds__(256) coarse_gemm_simt_kernel(int nq, int nc, int d, int dk, const float *__restrict__ Q,
                                                               const float *__restrict__ Cn,
                                                               const float *__restrict__ cnorm, int self_off,
                                                               float *__restrict__ grp_val,
                                                               uint32_t *__restrict__ grp_idx)
{
    extern __shared__ __align__(16) float sm[];
    float *As = sm;                      // [CG_BK][CG_BM]
    float *Bs = As + CG_BK * CG_BM;      // [CG_BK][CG_BN]
    float *D = Bs + CG_BK * CG_BN;       // [CG_BM][CG_BN + 1]
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int g = blockIdx.x, ngroups = gridDim.x;
    const int q0 = blockIdx.y * CG_BM, c0 = g * CG_BN;
    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; i++)
#pragma unroll
        for (int j = 0; j < 8; j++) acc[i][j] = 0.f;

    const int lrow = tid >> 1, lk = (tid & 1) * 4;
    for (int k0 = 0; k0 < dk; k0 += CG_BK) {
        float4 av = make_float4(0, 0, 0, 0), bv = make_float4(0, 0, 0, 0);
        if (q0 + lrow < nq && k0 + lk < dk) av = *reinterpret_cast<const float4 *>(Q + (size_t)(q0 + lrow) * d + k0 + lk);
        if (c0 + lrow < nc && k0 + lk < dk) bv = *reinterpret_cast<const float4 *>(Cn + (size_t)(c0 + lrow) * d + k0 + lk);
        As[(lk + 0) * CG_BM + lrow] = av.x; As[(lk + 1) * CG_BM + lrow] = av.y;
        As[(lk + 2) * CG_BM + lrow] = av.z; As[(lk + 3) * CG_BM + lrow] = av.w;
        Bs[(lk + 0) * CG_BN + lrow] = bv.x; Bs[(lk + 1) * CG_BN + lrow] = bv.y;
        Bs[(lk + 2) * CG_BN + lrow] = bv.z; Bs[(lk + 3) * CG_BN + lrow] = bv.w;
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < CG_BK; kk++) {
            float a[8], b[8];
            const float4 a0 = *reinterpret_cast<const float4 *>(As + kk * CG_BM + ty * 8);
            const float4 a1 = *reinterpret_cast<const float4 *>(As + kk * CG_BM + ty * 8 + 4);
            const float4 b0 = *reinterpret_cast<const float4 *>(Bs + kk * CG_BN + tx * 8);
            const float4 b1 = *reinterpret_cast<const float4 *>(Bs + kk * CG_BN + tx * 8 + 4);
            a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w; a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
            b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w; b[4] = b1.x; b[5] = b1.y; b[6] = b1.z; b[7] = b1.w;
#pragma unroll
            for (int i = 0; i < 8; i++)
#pragma unroll
                for (int j = 0; j < 8; j++) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const int c = c0 + tx * 8 + j;
        const float cn = c < nc ? cnorm[c] : 0.f;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const int q = q0 + ty * 8 + i;
            float v = cn - 2.f * acc[i][j];
            if (c >= nc || (self_off >= 0 && c == q + self_off)) v = __int_as_float(0x7f800000);
            D[(ty * 8 + i) * (CG_BN + 1) + tx * 8 + j] = v;
        }
    }
    __syncthreads();
    if (tid < CG_BM && q0 + tid < nq) {
        float v[COARSE_T];
        uint32_t ix[COARSE_T];
#pragma unroll
        for (int t = 0; t < COARSE_T; t++) { v[t] = __int_as_float(0x7f800000); ix[t] = 0xffffffffu; }
        const float *row = D + tid * (CG_BN + 1);
        for (int j = 0; j < CG_BN; j++) {
            float x = row[j];
            if (x < v[COARSE_T - 1]) {
                uint32_t xi = (uint32_t)(c0 + j);
#pragma unroll
                for (int t = 0; t < COARSE_T; t++) {
                    if (x < v[t]) {
                        const float tv = v[t]; const uint32_t ti = ix[t];
                        v[t] = x; ix[t] = xi; x = tv; xi = ti;
                    }
                }
            }
        }
        const size_t o = ((size_t)(q0 + tid) * ngroups + g) * COARSE_T;
#pragma unroll
        for (int t = 0; t < COARSE_T; t++) { grp_val[o + t] = v[t]; grp_idx[o + t] = ix[t]; }
    }
}

// ---------------------------------------------------------------------------------------------
// select + verify + exact re-rank.  One CTA per query.
// ---------------------------------------------------------------------------------------------
static __device__ __forceinline__ float key_to_float(uint32_t k)
{
    return __uint_as_float((k & 0x80000000u) ? (k ^ 0x80000000u) : ~k);   // inverse of fkey
}
static __device__ __forceinline__ uint32_t fkey(float f)
{
    const uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

#define SEL_MAX_FLAGGED 32

// SEL_CACHE: keys a thread keeps in registers across the radix passes (rows up to 256 * SEL_CACHE kept values)
template <int SEL_CACHE>
__global__ void __launch_bounds__(256) coarse_select_kernel(int nc, int d, int ngroups, int nprobe, int kprime,
                                                            int cand_cap, const float *__restrict__ Q,
                                                            const float *__restrict__ Cn,
                                                            const float *__restrict__ grp_val,
                                                            const uint32_t *__restrict__ grp_idx,
                                                            int packed, int kept, int self_off,
                                                            const uint32_t *__restrict__ cmax /* two-product GEMM: margin */,
                                                            const uint32_t *__restrict__ qstat /* packed == 2 */,
                                                            int gshift /* packed == 2: log2(tiles per record) */,
                                                            int pre /* packed == 2: 2 = a key stands for a column pair */,
                                                            int pcap /* pre == 2: capacity of the partner list */,
                                                            uint32_t idmul /* packed: column -> id multiplier, 0 = identity */,
                                                            uint32_t *__restrict__ err,
                                                            float *__restrict__ out_d, uint32_t *__restrict__ out_i)
{
    extern __shared__ __align__(16) uint8_t smraw[];
    float *sq = reinterpret_cast<float *>(smraw);                 // [d]
    uint32_t *cand = reinterpret_cast<uint32_t *>(sq + d);        // [cand_cap]
    float *cand_d = reinterpret_cast<float *>(cand + cand_cap);   // [cand_cap]
    uint32_t *pcand = reinterpret_cast<uint32_t *>(cand_d + cand_cap);   // [pcap] pre == 2: the other column of each pair
    __shared__ uint32_t hist[256];
    __shared__ uint32_t s_prefix, s_remaining, s_ncand, s_nflag, s_less, s_or, s_and, s_npc;
    __shared__ uint32_t s_flag[SEL_MAX_FLAGGED];
    const int q = blockIdx.x, tid = threadIdx.x;
    const int n = ngroups * COARSE_T;
    const float *gv = grp_val + (size_t)q * n;
    // packed (tensor-core GEMM): the centroid is group * COARSE_GROUP + low 7 bits of the value and
    // masked entries are >= 1e38; otherwise ids come from grp_idx and masked entries have id ~0
    // packed == 2 (impl 3): the entries are order-preserving integer keys (common.cuh), column in the low 7 bits,
    // padding columns / unused slots >= COARSE_KEY_INVALID_MIN
    const uint32_t *gi = packed ? nullptr : grp_idx + (size_t)q * n;
    const uint32_t *gk = reinterpret_cast<const uint32_t *>(gv);
    const uint32_t gsize = (uint32_t)COARSE_GROUP << gshift;          // centroids per record
    const uint32_t max_flag = SEL_MAX_FLAGGED >> gshift;              // expanded records (cand_cap holds 32 * 128 rows)
    const uint32_t kinv = packed == 2 ? COARSE_KEY_INVALID_MIN : fkey(1.0e38f);
    auto entry_key = [&](int i) -> uint32_t { return packed == 2 ? gk[i] : fkey(gv[i]); };
    auto is_self = [&](uint32_t id) -> bool { return self_off >= 0 && id == (uint32_t)(q + self_off); };
    // tensor-core GEMM: column -> centroid id (TcSide::idmul)
    auto col_id = [&](uint32_t col) -> uint32_t { return idmul ? (uint32_t)(((uint64_t)col * idmul) % (uint32_t)nc) : col; };
    auto entry_id = [&](int i) -> uint32_t {     // NOT filtered for the self match when packed (see the collect pass)
        if (!packed) return gi[i];
        uint32_t col;
        if (packed == 2) {
            if (gk[i] >= COARSE_KEY_INVALID_MIN) return 0xffffffffu;
            col = gk[i] & (gsize - 1u);      // super-groups: sub-tile in bits 7-8 = the column inside 512
        } else {
            const float v = gv[i];
            if (!(v < 1.0e38f)) return 0xffffffffu;
            col = __float_as_uint(v) & 127u;
        }
        return col_id((uint32_t)(i / COARSE_T) * gsize + col);
    };
    for (int i = tid; i < d; i += 256) sq[i] = Q[(size_t)q * d + i];
    if (tid == 0) { s_prefix = 0; s_remaining = (uint32_t)kprime; s_ncand = 0; s_nflag = 0; s_less = 0; s_or = 0u; s_and = 0xffffffffu; s_npc = 0; }
    __syncthreads();

    // ---- radix select: key of the kprime-th smallest kept value.  The keys a thread owns stay in registers
    // across the passes when the row is short enough (nc <= 65536); warp 0 scans the histogram.
    // Longer rows (nc > 2^17): the select runs on the 256 per-thread MINIMA instead.  Their kprime-th smallest is
    // an upper bound of the row's kprime-th smallest (they are 256 distinct entries of the row), and everything
    // below only needs tau >= the true value: the collect pass takes a few more candidates (expected
    // n * -ln(1 - kprime/256) / (n/256) instead of kprime), the exact re-rank sorts them out.  Two passes over the
    // row instead of five.
    const bool sampled = n > 256 * SEL_CACHE && kprime <= 192;
    const bool cached = sampled || n <= 256 * SEL_CACHE;
    const int nsel = sampled ? 512 : n;
    uint32_t kc[SEL_CACHE];
    if (sampled) {
        // the TWO smallest entries of every thread's share (512 distinct entries of the row): with kprime = 160 about
        // 169 row entries lie below their kprime-th smallest, against 251 with one minimum per thread
        uint32_t mk = 0xffffffffu, mk2 = 0xffffffffu;
        // entry i = 4 * group + rank-in-group: a thread that strode by 256 would only ever see ONE rank (the threads
        // with tid % 4 == 3 only 4th-best values) and the minima would say nothing about the row; rotating the
        // assignment by one entry per pass gives every thread all four ranks (loads stay coalesced)
        if (packed == 2) {
            // integer-key records are 16 bytes with ascending keys: a thread's two smallest entries are among the first
            // two keys of its records -- one 128-bit load and three min/max per record
            const uint4 *rec4 = reinterpret_cast<const uint4 *>(gk);
#pragma unroll 4
            for (int g = tid; g < ngroups; g += 256) {
                const uint4 r4 = rec4[g];
                mk2 = min(mk2, min(max(mk, r4.x), r4.y));
                mk = min(mk, r4.x);
            }
        } else {
            for (int j = 0; j * 256 < n; j++) {
                const int i = j * 256 + ((tid + j) & 255);
                if (i < n) {
                    const uint32_t kx = entry_key(i);
                    mk2 = min(mk2, max(mk, kx));
                    mk = min(mk, kx);
                }
            }
        }
#pragma unroll
        for (int j = 0; j < SEL_CACHE; j++) kc[j] = 0xffffffffu;
        kc[0] = mk;
        kc[1] = mk2;
    } else if (cached) {
#pragma unroll
        for (int j = 0; j < SEL_CACHE; j++) {
            const int i = tid + 256 * j;
            kc[j] = i < n ? entry_key(i) : 0xffffffffu;
        }
    }
    // bits every key shares need no pass (and would put all 2048 atomics of a pass on one counter): the keys of
    // one row usually share sign and exponent
    {
        // empty slots (>= 1e38: masked centroids, the unused 4th slot of the packed GEMM) take no part: they only
        // matter when kprime exceeds the number of valid entries, and then every valid entry is selected anyway
        uint32_t vor = 0u, vand = 0xffffffffu;
        if (cached) {
#pragma unroll
            for (int j = 0; j < SEL_CACHE; j++)
                if (tid + 256 * j < nsel && kc[j] < kinv) { vor |= kc[j]; vand &= kc[j]; }
        } else {
            for (int i = tid; i < n; i += 256) { const uint32_t kx = entry_key(i); if (kx < kinv) { vor |= kx; vand &= kx; } }
        }
        vor = __reduce_or_sync(IVF_FULL_MASK, vor);
        vand = __reduce_and_sync(IVF_FULL_MASK, vand);
        if ((tid & 31) == 0) { atomicOr(&s_or, vor); atomicAnd(&s_and, vand); }
    }
    __syncthreads();
    const uint32_t vdiff = s_or ^ s_and;
    int remaining_bits = vdiff ? 32 - __clz(vdiff) : 0;
    if (tid == 0) s_prefix = remaining_bits == 32 ? 0u : (s_and & ~((1u << remaining_bits) - 1u));
    __syncthreads();
    while (remaining_bits > 0) {
        const int w = remaining_bits < 8 ? remaining_bits : 8;
        const int shift = remaining_bits - w;
        const uint32_t dmask = (1u << w) - 1u;
        hist[tid] = 0;
        __syncthreads();
        const uint32_t prefix = s_prefix;
        const uint32_t pmask = remaining_bits == 32 ? 0u : (0xffffffffu << remaining_bits);
        if (cached) {
#pragma unroll
            for (int j = 0; j < SEL_CACHE; j++)
                if ((kc[j] & pmask) == prefix && tid + 256 * j < nsel) atomicAdd(&hist[(kc[j] >> shift) & dmask], 1u);
        } else {
            for (int i = tid; i < n; i += 256) {
                const uint32_t kx = entry_key(i);
                if ((kx & pmask) == prefix) atomicAdd(&hist[(kx >> shift) & dmask], 1u);
            }
        }
        __syncthreads();
        if (tid < 32) {
            // lane l owns bins 8l .. 8l+7
            uint32_t hb[8], mine = 0;
#pragma unroll
            for (int b = 0; b < 8; b++) { hb[b] = hist[tid * 8 + b]; mine += hb[b]; }
            uint32_t incl = mine;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(IVF_FULL_MASK, incl, o);
                if (tid >= o) incl += t;
            }
            const uint32_t rem = s_remaining;
            const uint32_t hit = __ballot_sync(IVF_FULL_MASK, incl >= rem);
            const int owner = hit ? __ffs(hit) - 1 : 31;   // kprime > n cannot happen (host clamps)
            if (tid == owner) {
                uint32_t cum = incl - mine;
                int b = 0;
#pragma unroll
                for (; b < 8; b++) {
                    if (cum + hb[b] >= rem) break;
                    cum += hb[b];
                }
                if (b == 8) b = 7;
                s_prefix = prefix | ((uint32_t)(tid * 8 + b) << shift);
                s_remaining = rem - cum;
                s_less += cum;
            }
        }
        __syncthreads();
        remaining_bits = shift;
    }
    uint32_t tau = s_prefix;       // key of the kprime-th smallest
    float pair_dmax = FLT_MAX;     // pre == 2: a partner column farther than this cannot be among the nprobe nearest
    if (packed == 2) {
        // one-product GEMM: each dropped cross term (q.c_lo always, q_lo.c when some query of the batch has a lo
        // part) moves a value by at most 2 * 2^-11 ||q|| ||c||; the true top-kprime all lie below the kprime-th
        // smallest approximate value + twice the total bound.  In key units (value * S) plus the quantisation of the
        // keys: one quantum of 128 for the rounding, one for the column bits.
        float qn = 0.f;
        const int dk = (d >> 4) ? ((d >> 4) << 4) : d;   // coarse_dims(d)
        for (int i = 0; i < dk; i++) qn = fmaf(sq[i], sq[i], qn);
        const float cmax2 = __uint_as_float(cmax[1]);
        const float S = coarse_key_scale(__uint_as_float(qstat[0]), cmax2);
        const float terms = qstat[1] ? 2.f : 1.f;
        const float eps = terms * 9.8e-4f * sqrtf(qn * cmax2);
        const float wide = 2.f * 1.02f * eps * S + 3.f * (float)gsize;
        if (tau < COARSE_KEY_INVALID_MIN) {
            const float room = (float)(COARSE_KEY_INVALID_MIN - 1u - tau);
            tau += (uint32_t)fminf(wide, room);
        }
        // pre == 2: the widened threshold as an exact distance.  key >> 7 counts quanta of 128 of t = v * S + 1.5 * 2^30
        // above 2^30 (v = ||c||^2 - 2 q.c, the distance without ||q||^2); every centroid of the true top-nprobe has
        // an approximate v <= v(tau) - eps, hence an exact distance <= v(tau) + ||q||^2.  Generous rounding head room:
        // a looser bound only lets a few more partner columns into the ranking.
        {
            const float tq = (float)(tau >> 7) * 128.f - COARSE_KEY_BIAS + 1024.f;
            const float vt = tq / S + qn;
            pair_dmax = vt + eps + fabsf(vt) * 1.0e-5f + fabsf(qn) * 1.0e-5f;
        }
    } else if (cmax) {
        // two-product GEMM: every kept value is within eps = 2 * 2^-11 ||q|| ||c|| of its three-product value, so the
        // true top-kprime all lie below (kprime-th smallest approximate value) + 2 eps; widen the threshold by that
        // (plus the 2^-16 relative perturbation of the packed column index, with some head room)
        float qn = 0.f;
        const int dk = (d >> 4) ? ((d >> 4) << 4) : d;   // coarse_dims(d)
        for (int i = 0; i < dk; i++) qn = fmaf(sq[i], sq[i], qn);
        const float tf = key_to_float(tau);
        const float eps = 9.8e-4f * sqrtf(qn * __uint_as_float(cmax[1]));
        const float wide = tf + 2.f * 1.02f * eps + fabsf(tf) * 6.2e-5f;
        if (tf < 1.0e38f) tau = fkey(wide);
    }
    auto push_id = [&](uint32_t id) {
        if (!is_self(id)) {
            const uint32_t p = atomicAdd(&s_ncand, 1u);
            if (p < (uint32_t)cand_cap) cand[p] = id;
        }
    };
    auto push_col = [&](uint32_t col) {      // integer keys: a GEMM column
        push_id(col_id(col));
        if (pre == 2) {      // the kept key is the smaller of columns (col, col ^ 1): the other one may pass too
            const uint32_t pc = col ^ 1u;
            if (pc < (uint32_t)nc) {
                const uint32_t pid = col_id(pc);
                if (!is_self(pid)) {
                    const uint32_t p = atomicAdd(&s_npc, 1u);
                    if (p < (uint32_t)pcap) pcand[p] = pid;
                }
            }
        }
    };
    const uint32_t tau_all = tau;
    if (packed == 2) {
        // ---- integer keys: ONE pass over the records (16 bytes each, keys ascending).  A record that may hide candidates
        // is flagged and expanded in full below; of the others every key <= tau is a candidate.
        const uint32_t tau = min(tau_all, COARSE_KEY_INVALID_MIN - 1u);     // an empty slot never passes
        for (int g = tid; g < ngroups; g += 256) {
            const uint4 r4 = reinterpret_cast<const uint4 *>(gk)[g];
            const uint32_t ev[4] = {r4.x, r4.y, r4.z, r4.w};
            if (ev[0] > tau) continue;
            bool hide;
            if (gshift) {
                // super-group record = the 4 smallest of what its tiles kept (3 each).  A tile may hide a value <= tau only
                // if all 3 of its kept values are <= tau: then either those 3 are in the record (same sub-tile bits) or
                // the record's 4th value is <= tau too.
                hide = ev[3] <= tau;
                if (!hide && ev[2] <= tau) {
                    const uint32_t s0 = (ev[0] >> 7) & 3u, s1 = (ev[1] >> 7) & 3u, s2 = (ev[2] >> 7) & 3u;
                    hide = (kept <= 3) && s0 == s1 && s1 == s2;
                    if (kept <= 2) hide = hide || s0 == s1 || s0 == s2 || s1 == s2;
                } else if (!hide && kept <= 2 && ev[1] <= tau) {
                    hide = ((ev[0] >> 7) & 3u) == ((ev[1] >> 7) & 3u);
                }
            } else {
                hide = (kept <= 2 ? ev[1] : (kept == 3 ? ev[2] : ev[3])) <= tau;
            }
            if (hide) {
                const uint32_t f = atomicAdd(&s_nflag, 1u);
                if (f < max_flag) s_flag[f] = (uint32_t)g;
                continue;
            }
#pragma unroll
            for (int j = 0; j < 4; j++)
                if (ev[j] <= tau) push_col((uint32_t)g * gsize + (ev[j] & (gsize - 1u)));
        }
        __syncthreads();
    }
    // ---- verify groups: a group whose COARSE_T-th kept value is <= tau may hide candidates
    for (int g = tid; packed != 2 && g < ngroups; g += 256) {
        const int last = g * COARSE_T + kept - 1;    // the group's last kept value (the GEMM keeps 3 or 4)
        const bool hide = entry_key(last) <= tau && (packed ? gv[last] < 1.0e38f : gi[last] != 0xffffffffu);
        if (hide) {
            const uint32_t f = atomicAdd(&s_nflag, 1u);
            if (f < max_flag) s_flag[f] = (uint32_t)g;
        }
    }
    __syncthreads();
    uint32_t nflag = s_nflag;
    if (nflag > max_flag) {
        if (tid == 0) atomicOr(err, 1u);
        nflag = max_flag;
    }
    // ---- collect: kept entries with key <= tau outside flagged groups, then flagged groups in full
    for (int i = tid; packed != 2 && i < n; i += 256) {
        const uint32_t kx = entry_key(i);
        if (kx > tau) continue;
        const uint32_t id = entry_id(i);
        if (id != 0xffffffffu) {
            const uint32_t g = (uint32_t)(i / COARSE_T);
            bool flagged = false;
            for (uint32_t f = 0; f < nflag; f++) flagged |= (s_flag[f] == g);
            if (!flagged) push_id(id);
        }
    }
    __syncthreads();
    for (uint32_t f = 0; f < nflag; f++) {
        const uint32_t g = s_flag[f];
        for (uint32_t j = tid; j < gsize; j += 256) {
            const uint32_t col = g * gsize + j;
            const uint32_t id = (packed && col < (uint32_t)nc) ? col_id(col) : col;
            if (col < (uint32_t)nc && !is_self(id)) {
                const uint32_t p = atomicAdd(&s_ncand, 1u);
                if (p < (uint32_t)cand_cap) cand[p] = id;
            }
        }
    }
    __syncthreads();
    uint32_t ncand = s_ncand;
    if (ncand > (uint32_t)cand_cap) {   // ties at tau can exceed the slack: flag, keep what fits
        if (tid == 0) atomicOr(err, 2u);
        ncand = cand_cap;
    }
    // ---- exact distances, reference numerics (8 lanes per candidate)
    const int lane = tid & 31, warp = tid >> 5, oct = lane >> 3, sub = lane & 7;
    for (uint32_t e0 = warp * 4; e0 < ncand; e0 += 32) {
        const uint32_t e = e0 + oct;
        const uint32_t id = e < ncand ? cand[e] : 0u;
        const float dist = l2sqr16_octet(sq, Cn + (size_t)id * d, d, sub);
        if (e < ncand && sub == 0) cand_d[e] = dist;
    }
    if (pre == 2) {
        // partner columns: exact distance, appended behind the candidates only when it can matter
        uint32_t npc = s_npc;
        if (npc > (uint32_t)pcap) {
            if (tid == 0) atomicOr(err, 2u);
            npc = pcap;
        }
        for (uint32_t e0 = warp * 4; e0 < npc; e0 += 32) {
            const uint32_t e = e0 + oct;
            const uint32_t id = e < npc ? pcand[e] : 0u;
            const float dist = l2sqr16_octet(sq, Cn + (size_t)id * d, d, sub);
            if (e < npc && sub == 0 && dist <= pair_dmax) {
                const uint32_t p = atomicAdd(&s_ncand, 1u);
                if (p < (uint32_t)cand_cap) { cand[p] = id; cand_d[p] = dist; }
            }
        }
        __syncthreads();
        ncand = s_ncand;
        if (ncand > (uint32_t)cand_cap) {
            if (tid == 0) atomicOr(err, 2u);
            ncand = cand_cap;
        }
    }
    __syncthreads();
    // ---- rank by (dist, idx) ascending == unload order of priority_queue<pair<float, idx_t>>
    for (uint32_t e = tid; e < ncand; e += 256) {
        const float ed = cand_d[e];
        const uint32_t ei = cand[e];
        uint32_t rank = 0;
        for (uint32_t o = 0; o < ncand; o++) rank += key_less(cand_d[o], cand[o], ed, ei) ? 1u : 0u;
        if (rank < (uint32_t)nprobe) {
            out_d[(size_t)q * nprobe + rank] = ed;
            out_i[(size_t)q * nprobe + rank] = ei;
        }
    }
    if (ncand < (uint32_t)nprobe) {
        if (tid == 0) atomicOr(err, 4u);
        for (uint32_t r = ncand + tid; r < (uint32_t)nprobe; r += 256) {
            out_d[(size_t)q * nprobe + r] = FLT_MAX;
            out_i[(size_t)q * nprobe + r] = 0;
        }
    }
}

size_t coarse_work_bytes(int nq, int nc, int kprime, int *cand_cap)
{
    const size_t ngroups = (nc + COARSE_GROUP - 1) / COARSE_GROUP;
    *cand_cap = kprime + 64 + COARSE_GROUP * SEL_MAX_FLAGGED;
    return (size_t)nq * ngroups * COARSE_T * 8 + 256;
}

cudaError_t launch_coarse_tc(cudaStream_t st, uint64_t *launches, int nq, int nc, int d, const float *queries,
                             const float *cnorm_gemm, const TcSide &S, float *grp_val, int nterms, int keep, int ikey,
                             int gshift, int pre);  // coarse_tc.cu

cudaError_t launch_coarse(cudaStream_t st, uint64_t *launches, int impl, int nq, int nc, int d, int nprobe,
                          int slack, const float *queries, const float *centroids, const float *cnorm_gemm,
                          const TcSide &S, const CoarseWork &W, int self_off, float *out_d, uint32_t *out_i,
                          cudaEvent_t ev_after_gemm)
{
    if (nq == 0) return cudaSuccess;
    const int ngroups = (nc + COARSE_GROUP - 1) / COARSE_GROUP;
    // values kept per group: the tensor-core epilogue keeps 3 instead of 4 when the groups are many (two min/max
    // fewer per accumulator element: -18 % GEMM time at 2^20 centroids); a group whose last kept value passes the
    // threshold is re-scanned in full by the select stage, which is rare only when the top-k spreads over many
    // groups (expected C(k,3) / ngroups^2 per query)
    // impl 3 keeps 2 when the groups are very many AND few neighbours are wanted (assign: nprobe 1): a group is expanded
    // by the select stage when its LAST kept value passes the threshold, expected C(k', kept) / ngroups^(kept-1) groups
    // per query -- at nprobe 128 / 2^20 centroids keeping 2 costs the select stage more than the GEMM gains (measured)
    const int kp = nprobe + slack;
    int kept = (impl >= 3 && ngroups >= 4096 && kp * kp < ngroups / 8) ? 2
                                                                        : ((impl >= 1 && ngroups >= 2048) ? 3 : COARSE_T);
    if (const char *e = getenv("IVFHNSW_B200_COARSE_KEEP")) {      // profiling switch
        const int kv = atoi(e);
        if (impl >= 1 && kv >= (impl >= 3 ? 2 : 3) && kv <= 4) kept = kv;
    }
    // impl 3 with many groups: the kept keys of 4 consecutive tiles leave the GEMM as ONE record of 4 (a 512-centroid
    // super-group): the select stage is bound by reading these records (three passes over 128 KB per query at 2^20
    // centroids); 3 kept per tile are needed for the hide test of the merged record
    int gshift = (impl >= 3 && ngroups >= 2048) ? 2 : 0;
    if (const char *e2 = getenv("IVFHNSW_B200_COARSE_GSHIFT")) gshift = (impl >= 3 && atoi(e2) == 2) ? 2 : 0;   // profiling switch
    // impl 3 with many groups: the epilogue reduces every column pair to its smaller key before the sorted insert (the
    // ALU pipe bounds that kernel); the select stage scores the other column of each surviving pair exactly
    int pre = (impl >= 3 && ngroups >= 2048 && kept <= 3) ? 2 : 1;
    if (const char *e3 = getenv("IVFHNSW_B200_COARSE_PRE")) pre = (impl >= 3 && kept <= 3 && atoi(e3) == 2) ? 2 : 1;   // profiling switch
    cudaError_t e;
    if (impl >= 1) {
        e = launch_coarse_tc(st, launches, nq, nc, d, queries, S.cnp, S, W.grp_val, impl >= 2 ? 1 : 2, kept,
                             impl >= 3 ? 1 : 0, gshift, pre);
        if (e != cudaSuccess) return e;
    } else {
        const size_t sm = (size_t)(CG_BK * CG_BM + CG_BK * CG_BN + CG_BM * (CG_BN + 1)) * sizeof(float);
        e = cudaFuncSetAttribute(coarse_gemm_simt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        if (e != cudaSuccess) return e;
        dim3 grid(ngroups, (nq + CG_BM - 1) / CG_BM);
        coarse_gemm_simt_kernel<<<grid, 256, sm, st>>>(nq, nc, d, coarse_dims(d), queries, centroids, cnorm_gemm, self_off,
                                                      W.grp_val, W.grp_idx);
        ++*launches;
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    if (ev_after_gemm) {
        e = cudaEventRecord(ev_after_gemm, st);
        if (e != cudaSuccess) return e;
    }
    const int nrec = gshift ? (ngroups + 3) >> 2 : ngroups;      // records per query the select stage reads
    int avail = nc - (self_off >= 0 ? 1 : 0);
    int kprime = nprobe + slack + ((impl >= 1 && self_off >= 0) ? 1 : 0);   // the packed GEMM keeps the self match
    if (kprime > avail) kprime = avail;
    if (kprime > nrec * (gshift ? COARSE_T : kept)) kprime = nrec * (gshift ? COARSE_T : kept);
    if (kprime < 1) kprime = 1;
    // partner columns of the pair prefilter: one per candidate outside the expanded groups (candidates: kprime, plus the
    // overshoot of the sampled threshold and the margin band)
    const int pcap = pre == 2 ? kprime + 64 + 512 : 0;
    const size_t sm2 = (size_t)d * 4 + (size_t)W.cand_cap * 8 + (size_t)pcap * 4;
    if (nrec * COARSE_T <= 256 * 8) {
        e = cudaFuncSetAttribute(coarse_select_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm2);
        if (e != cudaSuccess) return e;
        coarse_select_kernel<8><<<nq, 256, sm2, st>>>(nc, d, nrec, nprobe, kprime, W.cand_cap, queries, centroids,
                                                      W.grp_val, W.grp_idx, impl >= 3 ? 2 : (impl >= 1 ? 1 : 0), kept, self_off, impl >= 2 ? S.cmax : nullptr, S.qstat, gshift, pre, pcap, impl >= 1 ? S.idmul : 0u, W.err, out_d, out_i);
    } else {
        e = cudaFuncSetAttribute(coarse_select_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm2);
        if (e != cudaSuccess) return e;
        coarse_select_kernel<16><<<nq, 256, sm2, st>>>(nc, d, nrec, nprobe, kprime, W.cand_cap, queries, centroids,
                                                       W.grp_val, W.grp_idx, impl >= 3 ? 2 : (impl >= 1 ? 1 : 0), kept, self_off, impl >= 2 ? S.cmax : nullptr, S.qstat, gshift, pre, pcap, impl >= 1 ? S.idmul : 0u, W.err, out_d, out_i);
    }
    ++*launches;
    return cudaGetLastError();
}

}  // namespace ivfb200
