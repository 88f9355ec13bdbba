// C ABI of libivfhnsw_b200.so (declared in include/ivfhnsw_b200.h): handle state, host-side
// orchestration of the kernels in scan.cu / coarse.cu / build.cu / dense.cu.  No CPU fallback:
// every compute entry point needs a CUDA device and fails loudly otherwise.
#include <cuda_runtime.h>
#include <cub/cub.cuh>

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "common.cuh"
#include "handle.h"

namespace ivfb200 {
cudaError_t launch_group_points(cudaStream_t st, uint64_t *launches, int mode, int nc, int d, int nsubc,
                                const float *x, const uint32_t *order, const uint32_t *grp_off,
                                const float *centroids, const uint32_t *nn_idx, const float *icd,
                                const float *alphas, float *num, float *den, uint8_t *sub);
cudaError_t launch_slot_scan(cudaStream_t st, uint64_t *launches, int nc, uint32_t align, const uint32_t *size,
                             uint32_t *slot, uint32_t *raw_off, uint32_t *total_slots, int keep_b, int keep_e);
cudaError_t launch_scatter_lists2(cudaStream_t st, uint64_t *launches, size_t n, int M, int key_div,
                                  const uint32_t *order, const uint32_t *sorted_key, const uint32_t *raw_off,
                                  const uint32_t *list_slot, const uint32_t *ids, const uint8_t *codes,
                                  const uint8_t *norm_codes, uint8_t *out_codes, uint8_t *out_norm, uint8_t *out_sub,
                                  uint32_t *out_ids, uint32_t keep_b, uint32_t keep_e);
cudaError_t launch_gather_lists(cudaStream_t st, uint64_t *launches, int nc, int M, const uint32_t *list_size,
                                const uint32_t *list_slot, const uint32_t *raw_off, const uint8_t *codes,
                                const uint8_t *norm_codes, const uint32_t *ids, uint8_t *out_codes, uint8_t *out_norm,
                                uint32_t *out_ids);
}  // namespace ivfb200

using namespace ivfb200;

namespace {

__global__ void iota_kernel(size_t n, uint32_t *out)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        out[i] = (uint32_t)i;
}
__global__ void make_key_kernel(size_t n, const uint32_t *list, const uint8_t *sub, uint32_t mul, uint32_t *key)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        key[i] = list[i] * mul + (sub ? (uint32_t)sub[i] : 0u);
}
__global__ void key_to_list_kernel(size_t n, const uint32_t *key, uint32_t div, uint32_t *list)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        list[i] = key[i] / div;
}
// expand CSR offsets into one list id (and sub-group id) per record
__global__ void csr_to_keys_kernel(int nc, int nsubc, const int64_t *list_off, const uint32_t *sub_sizes,
                                   uint32_t *list, uint8_t *sub)
{
    const int c = blockIdx.x;
    const int64_t beg = list_off[c], end = list_off[c + 1];
    for (int64_t j = beg + threadIdx.x; j < end; j += blockDim.x) list[j] = (uint32_t)c;
    if (sub && threadIdx.x == 0) {
        int64_t pos = beg;
        for (int s = 0; s < nsubc; s++) {
            const uint32_t n = sub_sizes[(size_t)c * nsubc + s];
            for (uint32_t j = 0; j < n && pos < end; j++) sub[pos++] = (uint8_t)s;
        }
    }
}
__global__ void reduce_u32_kernel(int n, const uint32_t *a, const uint32_t *b, unsigned long long *out)
{
    unsigned long long sa = 0, sb = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) { sa += a[i]; sb += b[i]; }
    for (int off = 16; off > 0; off >>= 1) {
        sa += __shfl_xor_sync(0xffffffffu, sa, off);
        sb += __shfl_xor_sync(0xffffffffu, sb, off);
    }
    if ((threadIdx.x & 31) == 0) { atomicAdd(&out[0], sa); atomicAdd(&out[1], sb); }
}

__global__ void fill_u32_kernel(size_t n, uint32_t v, uint32_t *out)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) out[i] = v;
}
// a group may be encoded once: seen[g] accumulates the points of earlier flushes; flag = first offending group + 1
__global__ void group_seen_kernel(int nc, const uint32_t *gsize, uint32_t *seen, uint32_t *flag)
{
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= nc || gsize[g] == 0) return;
    if (seen[g]) atomicMax(flag, (uint32_t)g + 1u);
    seen[g] += gsize[g];
}
// out[0] = max(out[0], max_i key[i])
__global__ void max_u32_kernel(size_t n, const uint32_t *key, uint32_t *out)
{
    uint32_t m = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) m = max(m, key[i]);
    m = __reduce_max_sync(0xffffffffu, m);
    if ((threadIdx.x & 31) == 0) atomicMax(out, m);
}

unsigned grid_for(size_t n, int per = 256, unsigned cap = 148 * 16)
{
    size_t b = (n + per - 1) / per;
    if (b < 1) b = 1;
    return (unsigned)std::min<size_t>(b, cap);
}

}  // namespace

static int bind(ivfhnsw_b200_t *h)
{
    CU(cudaSetDevice(h->device));
    return 0;
}

// The coarse select stage records overflows (more flagged groups / candidates than its tables hold) in a sticky
// device word instead of returning an incomplete probe list silently.  Every entry point that ends in a stream
// synchronisation reads the word back (4 bytes) and fails loudly; the asynchronous *_dev entry points are covered by
// the next ivfhnsw_b200_sync() / host-pointer call on the handle.  Synchronises the stream.
static int check_coarse_flags(ivfhnsw_b200_t *h)
{
    unsigned long long w = 0;
    CU(cudaMemcpyAsync(&w, h->counters.as<unsigned long long>() + 4, 8, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    if (w & 0xffffffffull) {
        CU(cudaMemsetAsync(h->counters.as<unsigned long long>() + 4, 0, 8, h->stream));
        CU(cudaStreamSynchronize(h->stream));
        return h->fail(IVFHNSW_B200_ECUDA,
                       "coarse select overflow (flags 0x%llx: 1 = too many 128-centroid groups had to be expanded, 2 = too many "
                       "candidates tie at the threshold, 4 = fewer candidates than nprobe): the probe lists of this call are "
                       "incomplete.  nprobe is too large for nc (keep nprobe + coarse_slack below ~nc/128) or the centroid "
                       "table has many duplicates", w & 0xffffffffull);
    }
    return 0;
}

// list ids handed in by the caller (precomputed_idx files, add_group) must address an existing list
static int check_list_ids_dev(ivfhnsw_b200_t *h, size_t n, const uint32_t *idx_dev, const char *what)
{
    if (n == 0) return 0;
    CU(h->w_small.ensure(64));
    CU(cudaMemsetAsync(h->w_small.p, 0, 4, h->stream));
    max_u32_kernel<<<grid_for(n), 256, 0, h->stream>>>(n, idx_dev, h->w_small.as<uint32_t>());
    ++h->stats.kernel_launches;
    uint32_t m = 0;
    CU(cudaMemcpyAsync(&m, h->w_small.p, 4, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    if (m >= h->nc)
        return h->fail(IVFHNSW_B200_EINVAL, "%s: list id %u out of range (ncentroids = %zu): stale precomputed_idxs file?", what, m, h->nc);
    return 0;
}

extern "C" {

int ivfhnsw_b200_abi_version(void) { return IVFHNSW_B200_ABI; }

int ivfhnsw_b200_create(ivfhnsw_b200_t **out, size_t d, size_t nc, size_t code_size, size_t nbits, size_t nsubc,
                        int device)
{
    if (!out) return IVFHNSW_B200_EINVAL;
    *out = nullptr;
    ivfhnsw_b200_t *h = new ivfhnsw_b200();
    *out = h;  // returned even on failure so that last_error() is readable; caller destroys it
    h->device = device;
    if (const char *e = getenv("IVFHNSW_B200_COARSE_IMPL")) h->coarse_impl = atol(e);   // A/B switch for profiling
    h->d = d; h->nc = nc; h->M = code_size; h->nsubc = nsubc;
    h->c_begin = 0; h->c_end = (uint32_t)nc;
    if (nbits != 8) return h->fail(IVFHNSW_B200_EINVAL, "nbits_per_idx must be 8 (got %zu)", nbits);
    if (!(code_size == 8 || code_size == 16 || code_size == 32 || code_size == 64))
        return h->fail(IVFHNSW_B200_EINVAL, "bytes_per_code must be 8, 16, 32 or 64 (got %zu)", code_size);
    if (d == 0 || d % code_size != 0 || d % 4 != 0)
        return h->fail(IVFHNSW_B200_EINVAL, "dim %zu must be a multiple of 4 and of bytes_per_code %zu", d, code_size);
    if (nc == 0 || nc >= (1ull << 31)) return h->fail(IVFHNSW_B200_EINVAL, "bad ncentroids %zu", nc);
    if (nsubc > 256) return h->fail(IVFHNSW_B200_EINVAL, "nsubcentroids %zu > 256 unsupported", nsubc);
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0 || device < 0 || device >= ndev)
        return h->fail(IVFHNSW_B200_ENODEV, "no usable CUDA device %d (%s); this library has no CPU fallback", device,
                       e == cudaSuccess ? "device index out of range" : cudaGetErrorString(e));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10)
        return h->fail(IVFHNSW_B200_ENODEV, "device %d is sm_%d%d; kernels are built for sm_100a only", device, prop.major,
                       prop.minor);
    CU(cudaSetDevice(device));
    CU(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    for (int i = 0; i < 10; i++) CU(cudaEventCreate(&h->ev[i]));
    CU(h->counters.ensure(64));
    CU(cudaMemsetAsync(h->counters.p, 0, 64, h->stream));
    CU(h->centroid_norms.ensure(nc * 4));
    CU(cudaMemsetAsync(h->centroid_norms.p, 0, nc * 4, h->stream));
    if (nsubc) {
        CU(h->alphas.ensure(nc * 4));
        CU(cudaMemsetAsync(h->alphas.p, 0, nc * 4, h->stream));
    }
    CU(cudaStreamSynchronize(h->stream));
    return 0;
}

void ivfhnsw_b200_destroy(ivfhnsw_b200_t *h)
{
    if (!h) return;
    if (h->stream) {
        cudaSetDevice(h->device);
        cudaStreamSynchronize(h->stream);
    }
    DevBuf *bufs[] = {&h->centroids, &h->centroid_norms, &h->cnorm_gemm, &h->cent_f16, &h->pqc, &h->norm_table,
                      &h->opqA, &h->list_size, &h->list_slot, &h->raw_off, &h->codes, &h->norm_codes, &h->sub_ids,
                      &h->ids, &h->nn_idx, &h->sub_sizes, &h->alphas, &h->icd, &h->st_list, &h->st_ids, &h->st_codes,
                      &h->st_norm, &h->st_sub, &h->st_x, &h->rs_list, &h->rs_ids, &h->grp_seen, &h->w_x, &h->w_xrot, &h->w_cd, &h->w_ci, &h->w_seg,
                      &h->w_sub_base, &h->w_qsd, &h->w_g1, &h->w_part, &h->w_lut, &h->w_coarse, &h->w_xf16, &h->w_out_d, &h->w_out_l,
                      &h->w_a, &h->w_b, &h->w_c, &h->w_d2, &h->w_e, &h->w_f, &h->w_cub, &h->w_small, &h->counters};
    for (DevBuf *b : bufs) b->release();
    for (int i = 0; i < 10; i++)
        if (h->ev[i]) cudaEventDestroy(h->ev[i]);
    if (h->stream && h->own_stream) cudaStreamDestroy(h->stream);
    delete h;
}

const char *ivfhnsw_b200_last_error(const ivfhnsw_b200_t *h) { return h ? h->err.c_str() : "null handle"; }

int ivfhnsw_b200_set_shard(ivfhnsw_b200_t *h, size_t c_begin, size_t c_end)
{
    if (c_begin > c_end || c_end > h->nc) return h->fail(IVFHNSW_B200_EINVAL, "bad shard range");
    if (h->have_lists && !h->st_dirty && (c_begin < h->res_begin || c_end > h->res_end))
        return h->fail(IVFHNSW_B200_ESTATE, "shard [%zu, %zu) is not resident: the lists were built for [%u, %u)", c_begin,
                       c_end, h->res_begin, h->res_end);
    h->c_begin = (uint32_t)c_begin;
    h->c_end = (uint32_t)c_end;
    return 0;
}

static int upload(ivfhnsw_b200_t *h, DevBuf &b, const void *src, size_t bytes)
{
    CU(b.ensure(bytes));
    CU(cudaMemcpyAsync(b.p, src, bytes, cudaMemcpyHostToDevice, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return 0;
}

int ivfhnsw_b200_set_centroids(ivfhnsw_b200_t *h, const float *c)
{
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    if (int r = upload(h, h->centroids, c, h->nc * h->d * 4)) return r;
    h->have_centroids = true;
    h->gemm_norms_valid = false;
    h->f16_valid = false;
    if (!h->have_lists) h->have_group_meta = false;
    return 0;
}
int ivfhnsw_b200_set_pq(ivfhnsw_b200_t *h, const float *p)
{
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    if (int r = upload(h, h->pqc, p, h->d * 256 * 4)) return r;
    h->have_pq = true;
    return 0;
}
int ivfhnsw_b200_set_norm_pq(ivfhnsw_b200_t *h, const float *t)
{
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    if (int r = upload(h, h->norm_table, t, 256 * 4)) return r;
    h->have_norm = true;
    return 0;
}
int ivfhnsw_b200_set_opq(ivfhnsw_b200_t *h, const float *A)
{
    if (!A) { h->have_opq = false; return 0; }
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    if (int r = upload(h, h->opqA, A, h->d * h->d * 4)) return r;
    h->have_opq = true;
    return 0;
}
int ivfhnsw_b200_set_params(ivfhnsw_b200_t *h, size_t nprobe, size_t max_codes, int do_opq, int do_pruning)
{
    if (nprobe == 0 || nprobe > h->nc) return h->fail(IVFHNSW_B200_EINVAL, "nprobe %zu out of range (nc = %zu)", nprobe, h->nc);
    if (nprobe > 2048) return h->fail(IVFHNSW_B200_EINVAL, "nprobe %zu > 2048 unsupported", nprobe);
    h->nprobe = nprobe;
    h->max_codes = max_codes;
    h->do_opq = do_opq != 0;
    h->do_pruning = do_pruning != 0;
    return 0;
}
int ivfhnsw_b200_set_option(ivfhnsw_b200_t *h, const char *name, long v)
{
    if (!strcmp(name, "coarse_slack")) h->coarse_slack = v;
    else if (!strcmp(name, "coarse_impl")) h->coarse_impl = v;
    else if (!strcmp(name, "scan_split")) h->scan_split = v;
    else if (!strcmp(name, "query_batch")) h->query_batch = v;
    else if (!strcmp(name, "debug")) h->debug = v;
    else if (!strcmp(name, "stage_bytes")) h->stage_bytes = v > 0 ? v : 1;
    else return h->fail(IVFHNSW_B200_EINVAL, "unknown option %s", name);
    return 0;
}
int ivfhnsw_b200_set_timing(ivfhnsw_b200_t *h, int t) { h->timing = t != 0; return 0; }
int ivfhnsw_b200_set_stream(ivfhnsw_b200_t *h, void *stream)
{
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    CU(cudaStreamSynchronize(h->stream));
    if (h->own_stream && h->stream) CU(cudaStreamDestroy(h->stream));
    if (stream) {
        h->stream = reinterpret_cast<cudaStream_t>(stream);
        h->own_stream = false;
    } else {
        CU(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
        h->own_stream = true;
    }
    return 0;
}
int ivfhnsw_b200_sync(ivfhnsw_b200_t *h)
{
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    return check_coarse_flags(h);   // synchronises the stream
}

int ivfhnsw_b200_compute_centroid_norms(ivfhnsw_b200_t *h)
{
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    if (!h->have_centroids) return h->fail(IVFHNSW_B200_ESTATE, "centroids not set");
    CU(launch_row_norms(h->stream, LAUNCHES, h->nc, (int)h->d, h->centroids.as<float>(), h->centroid_norms.as<float>()));
    CU(cudaStreamSynchronize(h->stream));
    h->have_cnorms = true;
    return 0;
}
int ivfhnsw_b200_get_centroid_norms(ivfhnsw_b200_t *h, float *out)
{
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    CU(cudaMemcpyAsync(out, h->centroid_norms.p, h->nc * 4, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return 0;
}
int ivfhnsw_b200_set_centroid_norms(ivfhnsw_b200_t *h, const float *in)
{
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    if (int r = upload(h, h->centroid_norms, in, h->nc * 4)) return r;
    h->have_cnorms = true;
    return 0;
}
int ivfhnsw_b200_get_centroids(ivfhnsw_b200_t *h, float *out)
{
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    if (!h->have_centroids) return h->fail(IVFHNSW_B200_ESTATE, "centroids not set");
    CU(cudaMemcpyAsync(out, h->centroids.p, h->nc * h->d * 4, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return 0;
}
int ivfhnsw_b200_rotate_quantizer(ivfhnsw_b200_t *h)
{
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    if (!h->have_opq) return h->fail(IVFHNSW_B200_ESTATE, "OPQ encoding is turned off (no matrix set)");
    if (!h->have_centroids) return h->fail(IVFHNSW_B200_ESTATE, "centroids not set");
    CU(h->w_a.ensure(h->nc * h->d * 4));
    CU(launch_rotate(h->stream, LAUNCHES, h->nc, (int)h->d, h->opqA.as<float>(), false, h->centroids.as<float>(),
                     h->w_a.as<float>()));
    CU(cudaMemcpyAsync(h->centroids.p, h->w_a.p, h->nc * h->d * 4, cudaMemcpyDeviceToDevice, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    h->gemm_norms_valid = false;
    h->f16_valid = false;
    return 0;
}

// ------------------------------------------------------------------------------------ coarse helpers
static bool use_tc(const ivfhnsw_b200_t *h) { return h->coarse_impl >= 1 && coarse_tc_supported((int)h->d); }

// carve the fp16 operand buffers: centroids [hi | lo | max bits], queries [hi | lo | fac]
static TcSide tc_side(ivfhnsw_b200_t *h, size_t qb)
{
    TcSide S;
    S.dp = (int)((h->d + 7) / 8 * 8);
    S.chi = h->cent_f16.as<__half>();
    S.clo = S.chi ? S.chi + h->nc * S.dp : nullptr;
    S.cmax = S.chi ? reinterpret_cast<uint32_t *>(S.clo + h->nc * S.dp) : nullptr;
    S.cnp = S.chi ? reinterpret_cast<float *>(S.cmax + 16) : nullptr;
    S.idmul = coarse_col_multiplier(h->nc);
    S.qhi = h->w_xf16.as<__half>();
    S.qlo = S.qhi ? S.qhi + qb * S.dp : nullptr;
    S.qfac = S.qhi ? reinterpret_cast<float *>(S.qlo + qb * S.dp) : nullptr;
    S.qstat = S.qhi ? reinterpret_cast<uint32_t *>(S.qfac + ((qb + 3) & ~(size_t)3)) : nullptr;
    return S;
}

static int ensure_gemm_side(ivfhnsw_b200_t *h)
{
    if (!h->gemm_norms_valid) {
        // ||c||^2 over the dims fstdistfunc uses, for the GEMM ranking value
        CU(h->cnorm_gemm.ensure(h->nc * 4));
        CU(launch_row_norms(h->stream, LAUNCHES, h->nc, (int)h->d, h->centroids.as<float>(), h->cnorm_gemm.as<float>(),
                            coarse_dims((int)h->d)));
        h->gemm_norms_valid = true;
    }
    if (use_tc(h) && !h->f16_valid) {
        TcSide S = tc_side(h, 0);
        CU(h->cent_f16.ensure(h->nc * S.dp * 4 + 64 + h->nc * 4));
        S = tc_side(h, 0);
        CU(launch_tc_prepare_centroids(h->stream, LAUNCHES, h->nc, (int)h->d, h->centroids.as<float>(),
                                       h->cnorm_gemm.as<float>(), S));
        h->f16_valid = true;
    }
    return 0;
}

// exact nprobe nearest centroids for nq device-resident (already rotated) queries
int ivfb200_coarse_dev(ivfhnsw_b200_t *h, size_t nq, const float *xq, size_t nprobe, int exclude_self, float *out_d,
                        uint32_t *out_i)
{
    if (!h->have_centroids) return h->fail(IVFHNSW_B200_ESTATE, "centroids not set (build_quantizer)");
    if (int r = ensure_gemm_side(h)) return r;
    const int slack = (int)h->coarse_slack;
    int cand_cap = 0;
    const size_t per_q = coarse_work_bytes(1, (int)h->nc, (int)nprobe + slack, &cand_cap);
    size_t qb = std::min<size_t>(nq, std::max<size_t>(128, ((size_t)4 << 30) / per_q));
    CU(h->w_coarse.ensure(qb * per_q + 256));
    const int impl = use_tc(h) ? (int)std::min<long>(h->coarse_impl, 3) : 0;
    if (impl) CU(h->w_xf16.ensure(qb * ((h->d + 7) / 8 * 8) * 4 + qb * 4 + 512));
    const TcSide S = tc_side(h, qb);
    const size_t ngroups = (h->nc + COARSE_GROUP - 1) / COARSE_GROUP;
    for (size_t q0 = 0; q0 < nq; q0 += qb) {
        const size_t nb = std::min(qb, nq - q0);
        CoarseWork W;
        W.grp_val = h->w_coarse.as<float>();
        W.grp_idx = reinterpret_cast<uint32_t *>(W.grp_val + qb * ngroups * COARSE_T);
        W.err = reinterpret_cast<uint32_t *>(h->counters.as<unsigned long long>() + 4);
        W.cand = nullptr; W.cand_d = nullptr;
        W.cand_cap = cand_cap;
        // exclude_self needs global row ids: only used with q0 == 0-based centroid batches
        const float *cents = h->centroids.as<float>();
        if (h->timing) CU(cudaEventRecord(h->ev[6], h->stream));
        CU(launch_coarse(h->stream, LAUNCHES, impl, (int)nb, (int)h->nc, (int)h->d, (int)nprobe, slack,
                         xq + q0 * h->d, cents, h->cnorm_gemm.as<float>(), S, W,
                         exclude_self ? (int)q0 : -1, out_d + q0 * nprobe, out_i + q0 * nprobe,
                         h->timing ? h->ev[7] : nullptr));
        if (h->timing) {
            CU(cudaStreamSynchronize(h->stream));
            float ms = 0;
            CU(cudaEventElapsedTime(&ms, h->ev[6], h->ev[7]));
            h->stats.gemm_ms += ms;
        }
    }
    h->stats.coarse_flops += (uint64_t)2 * h->nc * h->d * nq;
    return 0;
}

int ivfhnsw_b200_coarse_dev(ivfhnsw_b200_t *h, size_t nq, const float *x_dev, size_t nprobe, float *dists_dev,
                            uint32_t *idxs_dev)
{
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    if (nprobe == 0 || nprobe > h->nc) return h->fail(IVFHNSW_B200_EINVAL, "nprobe out of range");
    const float *xq = x_dev;
    if (h->do_opq) {
        if (!h->have_opq) return h->fail(IVFHNSW_B200_ESTATE, "do_opq set but no OPQ matrix");
        CU(h->w_xrot.ensure(nq * h->d * 4));
        CU(launch_rotate(h->stream, LAUNCHES, nq, (int)h->d, h->opqA.as<float>(), false, x_dev, h->w_xrot.as<float>()));
        xq = h->w_xrot.as<float>();
    }
    return ivfb200_coarse_dev(h, nq, xq, nprobe, 0, dists_dev, idxs_dev);
}

int ivfhnsw_b200_coarse(ivfhnsw_b200_t *h, size_t nq, const float *x, size_t nprobe, float *dists, uint32_t *idxs)
{
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    if (nq == 0) return 0;
    CU(h->w_x.ensure(nq * h->d * 4));
    CU(h->w_cd.ensure(nq * nprobe * 4));
    CU(h->w_ci.ensure(nq * nprobe * 4));
    CU(cudaMemcpyAsync(h->w_x.p, x, nq * h->d * 4, cudaMemcpyHostToDevice, h->stream));
    if (int r = ivfhnsw_b200_coarse_dev(h, nq, h->w_x.as<float>(), nprobe, h->w_cd.as<float>(), h->w_ci.as<uint32_t>()))
        return r;
    CU(cudaMemcpyAsync(dists, h->w_cd.p, nq * nprobe * 4, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaMemcpyAsync(idxs, h->w_ci.p, nq * nprobe * 4, cudaMemcpyDeviceToHost, h->stream));
    return check_coarse_flags(h);
}

// assign for device-resident vectors (k = 1): labels_dev[n]
int ivfb200_assign_dev(ivfhnsw_b200_t *h, size_t n, const float *x_dev, uint32_t *labels_dev)
{
    CU(h->w_cd.ensure(n * 4));
    return ivfb200_coarse_dev(h, n, x_dev, 1, 0, h->w_cd.as<float>(), labels_dev);
}

int ivfhnsw_b200_assign_dev(ivfhnsw_b200_t *h, size_t n, const float *x_dev, uint32_t *labels_dev)
{
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    if (n == 0) return 0;
    return ivfb200_assign_dev(h, n, x_dev, labels_dev);
}

int ivfhnsw_b200_assign(ivfhnsw_b200_t *h, size_t n, const float *x, uint32_t *labels, size_t k)
{
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    if (k == 0 || k > h->nc || k > 2048) return h->fail(IVFHNSW_B200_EINVAL, "assign: k = %zu out of range", k);
    if (k > 1) {
        // IndexIVF_HNSW.cpp:68-72 takes searchKnn(x, k).top(): the k-th nearest centroid.  Exact coarse search with
        // nprobe = k, column k - 1 (every caller of the reference passes 1; kept for the interface)
        const size_t chunk = 1 << 16;
        const size_t cb = std::min(n, chunk);
        CU(h->w_x.ensure(cb * h->d * 4));
        CU(h->w_cd.ensure(cb * k * 4));
        CU(h->w_ci.ensure(std::max<size_t>(cb * k * 4, h->w_ci.cap)));
        for (size_t i0 = 0; i0 < n; i0 += chunk) {
            const size_t nb = std::min(chunk, n - i0);
            CU(cudaMemcpyAsync(h->w_x.p, x + i0 * h->d, nb * h->d * 4, cudaMemcpyHostToDevice, h->stream));
            if (int r = ivfb200_coarse_dev(h, nb, h->w_x.as<float>(), k, 0, h->w_cd.as<float>(), h->w_ci.as<uint32_t>()))
                return r;
            CU(cudaMemcpy2DAsync(labels + i0, 4, h->w_ci.as<uint32_t>() + (k - 1), k * 4, 4, nb, cudaMemcpyDeviceToHost,
                                 h->stream));
            if (int r = check_coarse_flags(h)) return r;
        }
        return 0;
    }
    const size_t chunk = 1 << 18;
    CU(h->w_x.ensure(std::min(n, chunk) * h->d * 4));
    CU(h->w_ci.ensure(std::max<size_t>(std::min(n, chunk) * 4, h->w_ci.cap)));
    for (size_t i0 = 0; i0 < n; i0 += chunk) {
        const size_t nb = std::min(chunk, n - i0);
        CU(cudaMemcpyAsync(h->w_x.p, x + i0 * h->d, nb * h->d * 4, cudaMemcpyHostToDevice, h->stream));
        if (int r = ivfb200_assign_dev(h, nb, h->w_x.as<float>(), h->w_ci.as<uint32_t>())) return r;
        CU(cudaMemcpyAsync(labels + i0, h->w_ci.p, nb * 4, cudaMemcpyDeviceToHost, h->stream));
        if (int r = check_coarse_flags(h)) return r;
    }
    return 0;
}

int ivfhnsw_b200_assign_u8(ivfhnsw_b200_t *h, size_t n, const uint8_t *x, uint32_t *labels)
{
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    const size_t chunk = 1 << 18;
    CU(h->w_x.ensure(std::min(n, chunk) * h->d * 4));
    CU(h->w_b.ensure(std::min(n, chunk) * h->d));
    CU(h->w_ci.ensure(std::max<size_t>(std::min(n, chunk) * 4, h->w_ci.cap)));
    for (size_t i0 = 0; i0 < n; i0 += chunk) {
        const size_t nb = std::min(chunk, n - i0);
        CU(cudaMemcpyAsync(h->w_b.p, x + i0 * h->d, nb * h->d, cudaMemcpyHostToDevice, h->stream));
        CU(launch_u8_to_f32(h->stream, LAUNCHES, nb * h->d, h->w_b.as<uint8_t>(), h->w_x.as<float>()));
        if (int r = ivfb200_assign_dev(h, nb, h->w_x.as<float>(), h->w_ci.as<uint32_t>())) return r;
        CU(cudaMemcpyAsync(labels + i0, h->w_ci.p, nb * 4, cudaMemcpyDeviceToHost, h->stream));
        if (int r = check_coarse_flags(h)) return r;
    }
    return 0;
}

// ------------------------------------------------------------------------------------ encoding
// codes / norm codes for n device vectors against base rows given by (idx[, sub]); chunked scratch
int ivfb200_encode_dev(ivfhnsw_b200_t *h, size_t n, const float *x_dev, const uint32_t *idx_dev,
                        const uint8_t *sub_dev, const float *alphas_dev, uint8_t *codes_dev, uint8_t *norm_dev,
                        float *norms_out, float *residual_out)
{
    // residual_out != nullptr: only the (rotated) residuals are produced (training); otherwise codes + norm codes
    // (norm_dev may be nullptr when only norms_out is wanted)
    if (!residual_out && (!h->have_pq || (norm_dev && !h->have_norm)))
        return h->fail(IVFHNSW_B200_ESTATE, "PQ codebooks not set (train_pq / read)");
    if (h->do_opq && !h->have_opq) return h->fail(IVFHNSW_B200_ESTATE, "do_opq set but no OPQ matrix");
    const int d = (int)h->d, M = (int)h->M, nsubc = (int)h->nsubc;
    const size_t chunk = std::min<size_t>(n, 1 << 19);
    CU(h->w_a.ensure(chunk * d * 4));
    CU(h->w_c.ensure(chunk * d * 4));
    const float *cen = h->centroids.as<float>();
    const uint32_t *nn = sub_dev ? h->nn_idx.as<uint32_t>() : nullptr;
    const float *al = sub_dev ? alphas_dev : nullptr;
    for (size_t i0 = 0; i0 < n; i0 += chunk) {
        const size_t nb = std::min(chunk, n - i0);
        float *R = h->w_a.as<float>(), *D = h->w_c.as<float>();
        const uint8_t *sb = sub_dev ? sub_dev + i0 : nullptr;
        // residuals (IndexIVF_HNSW.cpp:85-94)
        CU(launch_residual(h->stream, LAUNCHES, nb, d, x_dev + i0 * d, idx_dev + i0, sb, cen, nn, al, nsubc, -1.f,
                           nullptr, R));
        if (h->do_opq) {
            CU(launch_rotate(h->stream, LAUNCHES, nb, d, h->opqA.as<float>(), false, R, D));
            std::swap(R, D);
        }
        if (residual_out) {
            CU(cudaMemcpyAsync(residual_out + i0 * d, R, nb * d * 4, cudaMemcpyDeviceToDevice, h->stream));
            continue;
        }
        // encode / decode (:96-109)
        CU(launch_pq_encode(h->stream, LAUNCHES, nb, d, M, h->pqc.as<float>(), R, codes_dev + i0 * M));
        CU(launch_pq_decode(h->stream, LAUNCHES, nb, d, M, h->pqc.as<float>(), codes_dev + i0 * M, D));
        if (h->do_opq) {
            CU(launch_rotate(h->stream, LAUNCHES, nb, d, h->opqA.as<float>(), true, D, R));
            std::swap(R, D);
        }
        // reconstruct, norm, norm code (:111-121)
        CU(launch_residual(h->stream, LAUNCHES, nb, d, nullptr, idx_dev + i0, sb, cen, nn, al, nsubc, 1.f, D, R));
        CU(launch_norm_encode(h->stream, LAUNCHES, nb, d, R, h->norm_table.as<float>(), norm_dev ? norm_dev + i0 : nullptr,
                              norms_out ? norms_out + i0 : nullptr));
    }
    return 0;
}

// encoded records (st_*): list, id, code, norm code [, sub-group]
static int stage_reserve(ivfhnsw_b200_t *h, size_t extra)
{
    const size_t n = h->st_n + extra;
    CU(h->st_list.grow(n * 4, h->st_n * 4, h->stream));
    CU(h->st_ids.grow(n * 4, h->st_n * 4, h->stream));
    CU(h->st_codes.grow(n * h->M, h->st_n * h->M, h->stream));
    CU(h->st_norm.grow(n, h->st_n, h->stream));
    if (h->nsubc) CU(h->st_sub.grow(n, h->st_n, h->stream));
    return 0;
}
// raw points of a Grouping index waiting for their group to be complete (rs_*): list, id, vector
static int raw_reserve(ivfhnsw_b200_t *h, size_t extra)
{
    const size_t n = h->rs_n + extra;
    CU(h->rs_list.grow(n * 4, h->rs_n * 4, h->stream));
    CU(h->rs_ids.grow(n * 4, h->rs_n * 4, h->stream));
    CU(h->st_x.grow(n * h->d * 4, h->rs_n * h->d * 4, h->stream));
    return 0;
}

static int flush_groups_locked(ivfhnsw_b200_t *h);

static int unstage_existing_lists(ivfhnsw_b200_t *h);

static int add_batch_impl(ivfhnsw_b200_t *h, size_t n, const float *x, const uint32_t *xids, const uint32_t *idx,
                          uint8_t *codes_out, uint8_t *norm_out);

int ivfhnsw_b200_add_batch(ivfhnsw_b200_t *h, size_t n, const float *x, const uint32_t *xids, const uint32_t *idx)
{
    return add_batch_impl(h, n, x, xids, idx, nullptr, nullptr);
}
int ivfhnsw_b200_add_batch_codes(ivfhnsw_b200_t *h, size_t n, const float *x, const uint32_t *xids, const uint32_t *idx,
                                 uint8_t *codes_out, uint8_t *norm_codes_out)
{
    return add_batch_impl(h, n, x, xids, idx, codes_out, norm_codes_out);
}

static int add_batch_impl(ivfhnsw_b200_t *h, size_t n, const float *x, const uint32_t *xids, const uint32_t *idx,
                          uint8_t *codes_out, uint8_t *norm_out)
{
    std::lock_guard<std::mutex> lk(h->mu);
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    if (h->nsubc) return h->fail(IVFHNSW_B200_EINVAL, "add_batch on a Grouping index: use add_group");
    if (!h->have_centroids) return h->fail(IVFHNSW_B200_ESTATE, "centroids not set");
    if (n == 0) return 0;
    if (h->have_lists && !h->st_dirty) { if (int r = unstage_existing_lists(h)) return r; }
    if (int r = stage_reserve(h, n)) return r;
    const size_t chunk = 1 << 18;
    CU(h->w_x.ensure(std::min(n, chunk) * h->d * 4));
    for (size_t i0 = 0; i0 < n; i0 += chunk) {
        const size_t nb = std::min(chunk, n - i0);
        uint32_t *dl = h->st_list.as<uint32_t>() + h->st_n;
        CU(cudaMemcpyAsync(h->w_x.p, x + i0 * h->d, nb * h->d * 4, cudaMemcpyHostToDevice, h->stream));
        CU(cudaMemcpyAsync(h->st_ids.as<uint32_t>() + h->st_n, xids + i0, nb * 4, cudaMemcpyHostToDevice, h->stream));
        if (idx) {
            CU(cudaMemcpyAsync(dl, idx + i0, nb * 4, cudaMemcpyHostToDevice, h->stream));
            if (int r = check_list_ids_dev(h, nb, dl, "add_batch")) return r;
        } else {
            if (int r = ivfb200_assign_dev(h, nb, h->w_x.as<float>(), dl)) return r;
            if (int r = check_coarse_flags(h)) return r;
        }
        if (int r = ivfb200_encode_dev(h, nb, h->w_x.as<float>(), dl, nullptr, nullptr, h->st_codes.as<uint8_t>() + h->st_n * h->M,
                               h->st_norm.as<uint8_t>() + h->st_n, nullptr, nullptr))
            return r;
        if (codes_out)
            CU(cudaMemcpyAsync(codes_out + i0 * h->M, h->st_codes.as<uint8_t>() + h->st_n * h->M, nb * h->M,
                               cudaMemcpyDeviceToHost, h->stream));
        if (norm_out)
            CU(cudaMemcpyAsync(norm_out + i0, h->st_norm.as<uint8_t>() + h->st_n, nb, cudaMemcpyDeviceToHost, h->stream));
        CU(cudaStreamSynchronize(h->stream));
        h->st_n += nb;
    }
    h->st_dirty = true;
    return 0;
}

int ivfhnsw_b200_add_points_dev(ivfhnsw_b200_t *h, size_t n, const float *x_dev, const uint32_t *ids_dev,
                                const uint32_t *idx_dev)
{
    std::lock_guard<std::mutex> lk(h->mu);
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    if (n == 0) return 0;
    if (h->have_lists && !h->st_dirty) {
        if (h->nsubc) return h->fail(IVFHNSW_B200_ESTATE, "Grouping index already finalized: incremental add unsupported");
        if (int r = unstage_existing_lists(h)) return r;
    }
    const bool points = h->nsubc != 0;
    if (points) { if (int r = raw_reserve(h, n)) return r; }
    else if (int r = stage_reserve(h, n)) return r;
    uint32_t *dl = points ? h->rs_list.as<uint32_t>() + h->rs_n : h->st_list.as<uint32_t>() + h->st_n;
    uint32_t *di = points ? h->rs_ids.as<uint32_t>() + h->rs_n : h->st_ids.as<uint32_t>() + h->st_n;
    if (idx_dev) {
        CU(cudaMemcpyAsync(dl, idx_dev, n * 4, cudaMemcpyDeviceToDevice, h->stream));
        if (int r = check_list_ids_dev(h, n, dl, "add_points_dev")) return r;
    } else {
        if (int r = ivfb200_assign_dev(h, n, x_dev, dl)) return r;
        if (int r = check_coarse_flags(h)) return r;
    }
    CU(cudaMemcpyAsync(di, ids_dev, n * 4, cudaMemcpyDeviceToDevice, h->stream));
    if (points) {
        CU(cudaMemcpyAsync(h->st_x.as<float>() + h->rs_n * h->d, x_dev, n * h->d * 4, cudaMemcpyDeviceToDevice, h->stream));
        h->rs_n += n;
    } else {
        if (int r = ivfb200_encode_dev(h, n, x_dev, dl, nullptr, nullptr, h->st_codes.as<uint8_t>() + h->st_n * h->M,
                               h->st_norm.as<uint8_t>() + h->st_n, nullptr, nullptr))
            return r;
        h->st_n += n;
    }
    h->st_dirty = true;
    return 0;
}

int ivfhnsw_b200_add_group(ivfhnsw_b200_t *h, size_t g, size_t n, const float *x, const uint32_t *ids)
{
    std::lock_guard<std::mutex> lk(h->mu);  // called concurrently from an OpenMP loop in the reference's tests
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    if (!h->nsubc) return h->fail(IVFHNSW_B200_EINVAL, "add_group on an IVFADC index");
    if (g >= h->nc) return h->fail(IVFHNSW_B200_EINVAL, "group index out of range");
    if (h->have_lists && !h->st_dirty)
        return h->fail(IVFHNSW_B200_ESTATE, "Grouping index already finalized: incremental add_group unsupported");
    h->st_dirty = true;  // even empty groups mark the index as "being built" (nn lists are computed at finalize)
    if (n == 0) return 0;
    if (int r = raw_reserve(h, n)) return r;
    fill_u32_kernel<<<grid_for(n), 256, 0, h->stream>>>(n, (uint32_t)g, h->rs_list.as<uint32_t>() + h->rs_n);
    ++h->stats.kernel_launches;
    // pageable host copies return once the source has been staged: the caller may free x / ids right after
    CU(cudaMemcpyAsync(h->rs_ids.as<uint32_t>() + h->rs_n, ids, n * 4, cudaMemcpyHostToDevice, h->stream));
    CU(cudaMemcpyAsync(h->st_x.as<float>() + h->rs_n * h->d, x, n * h->d * 4, cudaMemcpyHostToDevice, h->stream));
    {   // pinned sources are copied asynchronously: wait, the caller owns the buffers again when we return
        cudaPointerAttributes pa{};
        if (cudaPointerGetAttributes(&pa, x) != cudaSuccess || pa.type != cudaMemoryTypeUnregistered) {
            (void)cudaGetLastError();
            CU(cudaStreamSynchronize(h->stream));
        }
    }
    h->rs_n += n;
    // one add_group call delivers a COMPLETE group (IndexIVF_HNSW_Grouping.cpp:43-157 encodes it on the spot): the
    // staged groups can be encoded as soon as enough of them are waiting -- the bounded pass of
    // tests/test_ivfhnsw_grouping_sift1b.cpp:138-201, without holding the raw vectors of the whole base set
    if (h->rs_n * h->d * 4 >= (size_t)h->stage_bytes) return flush_groups_locked(h);
    return 0;
}

int ivfhnsw_b200_flush_groups(ivfhnsw_b200_t *h)
{
    std::lock_guard<std::mutex> lk(h->mu);
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    if (!h->nsubc) return 0;
    return flush_groups_locked(h);
}

// ------------------------------------------------------------------------------------ CSR build
// staged records (st_list[, st_sub], st_ids, st_codes, st_norm) -> slot-space lists
static int build_csr(ivfhnsw_b200_t *h)
{
    const size_t n = h->st_n;
    const int nc = (int)h->nc, M = (int)h->M, nsubc = (int)h->nsubc;
    const uint32_t key_div = nsubc ? (uint32_t)nsubc : 1u;
    if ((uint64_t)nc * key_div >= (1ull << 32)) return h->fail(IVFHNSW_B200_EINVAL, "nc * nsubc overflows 32 bits");
    if (n >= (1ull << 31))
        return h->fail(IVFHNSW_B200_EINVAL, "%zu records staged: one device holds at most 2^31 - 1 (shard the lists with set_shard)", n);
    cudaStream_t st = h->stream;
    CU(h->list_size.ensure((size_t)nc * 4));
    CU(h->list_slot.ensure((size_t)nc * 4));
    CU(h->raw_off.ensure((size_t)(nc + 1) * 4));
    CU(cudaMemsetAsync(h->list_size.p, 0, (size_t)nc * 4, st));
    if (nsubc) {
        CU(h->sub_sizes.ensure((size_t)nc * nsubc * 4));
        CU(cudaMemsetAsync(h->sub_sizes.p, 0, (size_t)nc * nsubc * 4, st));
    }
    uint32_t *keys_in = nullptr, *keys_out = nullptr, *ord_in = nullptr, *ord_out = nullptr;
    if (n) {
        CU(h->w_d2.ensure(n * 16));
        keys_in = h->w_d2.as<uint32_t>(); keys_out = keys_in + n; ord_in = keys_out + n; ord_out = ord_in + n;
        make_key_kernel<<<grid_for(n), 256, 0, st>>>(n, h->st_list.as<uint32_t>(), nsubc ? h->st_sub.as<uint8_t>() : nullptr,
                                                     key_div, keys_in);
        iota_kernel<<<grid_for(n), 256, 0, st>>>(n, ord_in);
        h->stats.kernel_launches += 2;
        size_t tmp = 0;
        int bits = 1;
        while (((uint64_t)nc * key_div) >> bits) bits++;
        CU(cub::DeviceRadixSort::SortPairs(nullptr, tmp, keys_in, keys_out, ord_in, ord_out, (int)n, 0, bits, st));
        CU(h->w_cub.ensure(tmp + 16));
        CU(cub::DeviceRadixSort::SortPairs(h->w_cub.p, tmp, keys_in, keys_out, ord_in, ord_out, (int)n, 0, bits, st));
        h->stats.kernel_launches += 4;
        CU(launch_hist(st, LAUNCHES, n, h->st_list.as<uint32_t>(), h->list_size.as<uint32_t>()));
        if (nsubc) CU(launch_hist(st, LAUNCHES, n, keys_in, h->sub_sizes.as<uint32_t>()));
    }
    const uint32_t align = M <= 16 ? 16u : 32u;
    CU(h->w_small.ensure(64));
    CU(launch_slot_scan(st, LAUNCHES, nc, align, h->list_size.as<uint32_t>(), h->list_slot.as<uint32_t>(),
                        h->raw_off.as<uint32_t>(), h->w_small.as<uint32_t>(), (int)h->c_begin, (int)h->c_end));
    uint32_t tot[2] = {0, 0};
    CU(cudaMemcpyAsync(tot, h->w_small.p, 8, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    const uint64_t total_slots = ((uint64_t)tot[1] << 32) | tot[0];
    if (total_slots + 64 >= (1ull << 32)) return h->fail(IVFHNSW_B200_EINVAL, "index too large for 32-bit slot ids");
    const size_t nslots = (size_t)total_slots + 64;
    CU(h->codes.ensure(nslots * M));
    CU(h->norm_codes.ensure(nslots));
    CU(h->ids.ensure(nslots * 4));
    CU(cudaMemsetAsync(h->codes.p, 0, nslots * M, st));
    CU(cudaMemsetAsync(h->norm_codes.p, 0, nslots, st));
    CU(cudaMemsetAsync(h->ids.p, 0, nslots * 4, st));
    if (nsubc) {
        CU(h->sub_ids.ensure(nslots));
        CU(cudaMemsetAsync(h->sub_ids.p, 0, nslots, st));
    }
    if (n)
        CU(launch_scatter_lists2(st, LAUNCHES, n, M, (int)key_div, ord_out, keys_out, h->raw_off.as<uint32_t>(),
                                 h->list_slot.as<uint32_t>(), h->st_ids.as<uint32_t>(), h->st_codes.as<uint8_t>(),
                                 h->st_norm.as<uint8_t>(), h->codes.as<uint8_t>(), h->norm_codes.as<uint8_t>(),
                                 nsubc ? h->sub_ids.as<uint8_t>() : nullptr, h->ids.as<uint32_t>(), h->c_begin, h->c_end));
    CU(cudaStreamSynchronize(st));
    h->nslots = nslots;
    h->ntotal = n;
    h->have_lists = true;
    h->res_begin = h->c_begin;   // only these lists are resident from here on
    h->res_end = h->c_end;
    return 0;
}

// pull the current CSR back into the staging arrays so that more records can be appended
static int unstage_existing_lists(ivfhnsw_b200_t *h)
{
    const size_t n = h->ntotal;
    h->st_n = 0;
    if (n == 0) return 0;
    if (h->res_begin != 0 || h->res_end != h->nc)
        return h->fail(IVFHNSW_B200_ESTATE, "cannot append to lists built for the shard [%u, %u) only", h->res_begin, h->res_end);
    if (int r = stage_reserve(h, n)) return r;
    CU(launch_gather_lists(h->stream, LAUNCHES, (int)h->nc, (int)h->M, h->list_size.as<uint32_t>(),
                           h->list_slot.as<uint32_t>(), h->raw_off.as<uint32_t>(), h->codes.as<uint8_t>(),
                           h->norm_codes.as<uint8_t>(), h->ids.as<uint32_t>(), h->st_codes.as<uint8_t>(),
                           h->st_norm.as<uint8_t>(), h->st_ids.as<uint32_t>()));
    // list id per record from raw offsets
    CU(h->w_e.ensure((h->nc + 1) * 8));
    std::vector<uint32_t> ro(h->nc + 1);
    CU(cudaMemcpyAsync(ro.data(), h->raw_off.p, (h->nc + 1) * 4, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    std::vector<int64_t> off(h->nc + 1);
    for (size_t i = 0; i <= h->nc; i++) off[i] = ro[i];
    CU(cudaMemcpyAsync(h->w_e.p, off.data(), (h->nc + 1) * 8, cudaMemcpyHostToDevice, h->stream));
    csr_to_keys_kernel<<<(unsigned)h->nc, 128, 0, h->stream>>>((int)h->nc, 0, h->w_e.as<int64_t>(), nullptr,
                                                              h->st_list.as<uint32_t>(), nullptr);
    ++h->stats.kernel_launches;
    CU(cudaStreamSynchronize(h->stream));
    h->st_n = n;
    return 0;
}

// K-G3: nsubc nearest other centroids of every centroid (IndexIVF_HNSW_Grouping.cpp:47-63), exact, and the
// distances to them (== compute_inter_centroid_dists, :620-631).  Cached until the centroids change.
int ivfb200_ensure_group_meta(ivfhnsw_b200_t *h)
{
    if (h->have_group_meta) return 0;
    const int nc = (int)h->nc, nsubc = (int)h->nsubc;
    if ((size_t)nsubc + 1 > h->nc) return h->fail(IVFHNSW_B200_EINVAL, "nsubc + 1 > nc");
    CU(h->nn_idx.ensure((size_t)nc * nsubc * 4));
    CU(h->icd.ensure((size_t)nc * nsubc * 4));
    if (int r = ivfb200_coarse_dev(h, nc, h->centroids.as<float>(), nsubc, 1, h->icd.as<float>(), h->nn_idx.as<uint32_t>()))
        return r;
    if (int r = check_coarse_flags(h)) return r;
    h->have_group_meta = true;
    return 0;
}

// compute_alpha (:691-728) and compute_subcentroid_idxs (:673-689) for n device points
int ivfb200_group_stage(ivfhnsw_b200_t *h, size_t n, const float *x_dev, const uint32_t *list_dev, float *alphas_out,
                         uint8_t *sub_out, int zero_alphas, uint32_t *seen /* per-group point counts of earlier calls, or null */)
{
    const int nc = (int)h->nc, d = (int)h->d, nsubc = (int)h->nsubc;
    cudaStream_t st = h->stream;
    if (int r = ivfb200_ensure_group_meta(h)) return r;
    if (zero_alphas) CU(cudaMemsetAsync(alphas_out, 0, (size_t)nc * 4, st));
    if (n == 0) return 0;
    if (n >= (1ull << 31)) return h->fail(IVFHNSW_B200_EINVAL, "%zu points in one Grouping batch (limit 2^31 - 1)", n);
    // sort points by group (stable) -> order, grp_off
    CU(h->w_d2.ensure(n * 16));
    uint32_t *keys_out = h->w_d2.as<uint32_t>(), *ord_in = keys_out + n, *ord_out = ord_in + n;
    iota_kernel<<<grid_for(n), 256, 0, st>>>(n, ord_in);
    ++h->stats.kernel_launches;
    size_t tmp = 0;
    int bits = 1;
    while (((uint64_t)nc) >> bits) bits++;
    CU(cub::DeviceRadixSort::SortPairs(nullptr, tmp, list_dev, keys_out, ord_in, ord_out, (int)n, 0, bits, st));
    CU(h->w_cub.ensure(tmp + 16));
    CU(cub::DeviceRadixSort::SortPairs(h->w_cub.p, tmp, list_dev, keys_out, ord_in, ord_out, (int)n, 0, bits, st));
    h->stats.kernel_launches += 4;
    CU(h->w_e.ensure((size_t)(nc + 1) * 12 + 64));
    uint32_t *gsize = h->w_e.as<uint32_t>(), *gslot = gsize + nc, *goff = gslot + nc;
    CU(cudaMemsetAsync(gsize, 0, (size_t)nc * 4, st));
    CU(launch_hist(st, LAUNCHES, n, list_dev, gsize));
    CU(h->w_small.ensure(64));
    CU(launch_slot_scan(st, LAUNCHES, nc, 1, gsize, gslot, goff, h->w_small.as<uint32_t>(), 0, nc));
    if (seen) {
        uint32_t *flag = h->w_small.as<uint32_t>() + 8;
        CU(cudaMemsetAsync(flag, 0, 4, st));
        group_seen_kernel<<<(nc + 255) / 256, 256, 0, st>>>(nc, gsize, seen, flag);
        ++h->stats.kernel_launches;
        uint32_t bad = 0;
        CU(cudaMemcpyAsync(&bad, flag, 4, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        if (bad)
            return h->fail(IVFHNSW_B200_ESTATE, "group %u received points after it had been encoded: every flush (flush_groups, or the "
                           "automatic one of add_group) must hold COMPLETE groups -- alpha and the sub-groups are functions of "
                           "all of a group's points (IndexIVF_HNSW_Grouping.cpp:43-157)", bad - 1);
    }
    CU(h->w_f.ensure(n * 8));
    float *num = h->w_f.as<float>(), *den = num + n;
    CU(launch_group_points(st, LAUNCHES, 0, nc, d, nsubc, x_dev, ord_out, goff, h->centroids.as<float>(),
                           h->nn_idx.as<uint32_t>(), h->icd.as<float>(), nullptr, num, den, nullptr));
    CU(launch_alpha_reduce(st, LAUNCHES, nc, goff, num, den, alphas_out));
    CU(launch_group_points(st, LAUNCHES, 1, nc, d, nsubc, x_dev, ord_out, goff, h->centroids.as<float>(),
                           h->nn_idx.as<uint32_t>(), h->icd.as<float>(), alphas_out, nullptr, nullptr, sub_out));
    return 0;
}

// Grouping: alpha, sub-group ids and codes of the raw points staged so far (complete groups), appended to the
// encoded records; the raw vectors are dropped.  add_group body, IndexIVF_HNSW_Grouping.cpp:43-157.
static int flush_groups_locked(ivfhnsw_b200_t *h)
{
    const size_t n = h->rs_n;
    if (n == 0) return 0;
    if (!h->have_centroids) return h->fail(IVFHNSW_B200_ESTATE, "centroids not set");
    if (int r = stage_reserve(h, n)) return r;
    CU(h->grp_seen.ensure(h->nc * 4));
    if (!h->grp_seen_valid) {
        CU(cudaMemsetAsync(h->grp_seen.p, 0, h->nc * 4, h->stream));
        h->grp_seen_valid = true;
    }
    uint8_t *sub = h->st_sub.as<uint8_t>() + h->st_n;
    if (int r = ivfb200_group_stage(h, n, h->st_x.as<float>(), h->rs_list.as<uint32_t>(), h->alphas.as<float>(), sub, 0,
                                    h->grp_seen.as<uint32_t>()))
        return r;
    if (int r = ivfb200_encode_dev(h, n, h->st_x.as<float>(), h->rs_list.as<uint32_t>(), sub, h->alphas.as<float>(),
                                   h->st_codes.as<uint8_t>() + h->st_n * h->M, h->st_norm.as<uint8_t>() + h->st_n, nullptr,
                                   nullptr))
        return r;
    CU(cudaMemcpyAsync(h->st_list.as<uint32_t>() + h->st_n, h->rs_list.p, n * 4, cudaMemcpyDeviceToDevice, h->stream));
    CU(cudaMemcpyAsync(h->st_ids.as<uint32_t>() + h->st_n, h->rs_ids.p, n * 4, cudaMemcpyDeviceToDevice, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    h->st_n += n;
    h->rs_n = 0;
    return 0;
}

static int finalize_grouping(ivfhnsw_b200_t *h)
{
    if (int r = ivfb200_ensure_group_meta(h)) return r;   // nn lists / icd exist even for an empty index
    if (int r = flush_groups_locked(h)) return r;
    h->st_x.release();  // raw vectors are not needed any more (and the CSR build wants the memory)
    h->rs_list.release();
    h->rs_ids.release();
    h->grp_seen.release();
    h->grp_seen_valid = false;
    h->w_f.release();
    CU(h->st_sub.ensure(1));
    return build_csr(h);
}

int ivfhnsw_b200_finalize(ivfhnsw_b200_t *h)
{
    std::lock_guard<std::mutex> lk(h->mu);
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    if (h->have_lists && !h->st_dirty) return 0;
    if (!h->have_centroids) return h->fail(IVFHNSW_B200_ESTATE, "centroids not set");
    int r = h->nsubc ? finalize_grouping(h) : build_csr(h);
    if (r) return r;
    h->st_dirty = false;
    // the staged records now live in the CSR: give the staging and sort buffers back (26 + 16 GB at 1B vectors);
    // a later add_batch re-stages from the lists (unstage_existing_lists)
    DevBuf *tmp[] = {&h->st_list, &h->st_ids, &h->st_codes, &h->st_norm, &h->st_sub, &h->w_d2, &h->w_cub};
    for (DevBuf *b : tmp) b->release();
    h->st_n = 0;
    return 0;
}

int ivfhnsw_b200_load_lists(ivfhnsw_b200_t *h, const int64_t *list_off, const uint32_t *ids, const uint8_t *codes,
                            const uint8_t *norm_codes, const uint32_t *nn_idx, const uint32_t *sub_sizes,
                            const float *alphas, const float *icd)
{
    std::lock_guard<std::mutex> lk(h->mu);
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    const size_t nc = h->nc, nsubc = h->nsubc, M = h->M;
    if (list_off[0] != 0) return h->fail(IVFHNSW_B200_EINVAL, "list_off[0] must be 0");
    const size_t n = (size_t)list_off[nc];
    if (nsubc && (!nn_idx || !sub_sizes || !alphas || !icd))
        return h->fail(IVFHNSW_B200_EINVAL, "Grouping index needs nn_idx, sub_sizes, alphas, icd");
    h->st_n = 0;
    if (int r = stage_reserve(h, std::max<size_t>(n, 1))) return r;
    cudaStream_t st = h->stream;
    CU(h->w_e.ensure((nc + 1) * 8));
    CU(cudaMemcpyAsync(h->w_e.p, list_off, (nc + 1) * 8, cudaMemcpyHostToDevice, st));
    if (n) {
        CU(cudaMemcpyAsync(h->st_ids.p, ids, n * 4, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(h->st_codes.p, codes, n * M, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(h->st_norm.p, norm_codes, n, cudaMemcpyHostToDevice, st));
    }
    if (nsubc) {
        CU(h->nn_idx.ensure(nc * nsubc * 4));
        CU(h->icd.ensure(nc * nsubc * 4));
        CU(h->w_f.ensure(nc * nsubc * 4));
        CU(cudaMemcpyAsync(h->nn_idx.p, nn_idx, nc * nsubc * 4, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(h->icd.p, icd, nc * nsubc * 4, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(h->alphas.p, alphas, nc * 4, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(h->w_f.p, sub_sizes, nc * nsubc * 4, cudaMemcpyHostToDevice, st));
        h->have_group_meta = true;
    }
    csr_to_keys_kernel<<<(unsigned)nc, 128, 0, st>>>((int)nc, (int)nsubc, h->w_e.as<int64_t>(),
                                                    nsubc ? h->w_f.as<uint32_t>() : nullptr, h->st_list.as<uint32_t>(),
                                                    nsubc ? h->st_sub.as<uint8_t>() : nullptr);
    ++h->stats.kernel_launches;
    CU(cudaStreamSynchronize(st));
    h->st_n = n;
    if (int r = build_csr(h)) return r;
    h->st_dirty = false;
    DevBuf *tmp[] = {&h->st_list, &h->st_ids, &h->st_codes, &h->st_norm, &h->st_sub, &h->w_d2, &h->w_cub};
    for (DevBuf *b : tmp) b->release();
    h->st_n = 0;
    return 0;
}

int ivfhnsw_b200_list_sizes(ivfhnsw_b200_t *h, uint32_t *sizes)
{
    if (int r = ivfhnsw_b200_finalize(h)) return r;
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    if (!h->have_lists) return h->fail(IVFHNSW_B200_ESTATE, "no lists");
    CU(cudaMemcpyAsync(sizes, h->list_size.p, h->nc * 4, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return 0;
}

// Sharded construction: every rank ingests only the points of the lists it owns, so its per-list metadata is
// zero elsewhere; the ranks sum the arrays (NCCL all-reduce over disjoint supports) and install the result.
int ivfhnsw_b200_get_list_meta_dev(ivfhnsw_b200_t *h, uint32_t *list_size_dev, uint32_t *sub_sizes_dev, float *alphas_dev)
{
    if (int r = ivfhnsw_b200_finalize(h)) return r;
    std::lock_guard<std::mutex> lk(h->mu);
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    if (!h->have_lists) return h->fail(IVFHNSW_B200_ESTATE, "no lists");
    cudaStream_t st = h->stream;
    const size_t nc = h->nc, nsubc = h->nsubc;
    if (list_size_dev) CU(cudaMemcpyAsync(list_size_dev, h->list_size.p, nc * 4, cudaMemcpyDeviceToDevice, st));
    if (nsubc && sub_sizes_dev) CU(cudaMemcpyAsync(sub_sizes_dev, h->sub_sizes.p, nc * nsubc * 4, cudaMemcpyDeviceToDevice, st));
    if (nsubc && alphas_dev) CU(cudaMemcpyAsync(alphas_dev, h->alphas.p, nc * 4, cudaMemcpyDeviceToDevice, st));
    return 0;
}

int ivfhnsw_b200_set_list_meta_dev(ivfhnsw_b200_t *h, const uint32_t *list_size_dev, const uint32_t *sub_sizes_dev,
                                   const float *alphas_dev)
{
    if (int r = ivfhnsw_b200_finalize(h)) return r;
    std::lock_guard<std::mutex> lk(h->mu);
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    if (!h->have_lists) return h->fail(IVFHNSW_B200_ESTATE, "no lists");
    if (!list_size_dev || (h->nsubc && (!sub_sizes_dev || !alphas_dev)))
        return h->fail(IVFHNSW_B200_EINVAL, "set_list_meta: null array");
    cudaStream_t st = h->stream;
    const size_t nc = h->nc, nsubc = h->nsubc;
    // the resident lists must keep their sizes: their slots were laid out from them
    std::vector<uint32_t> mine(nc), all(nc);
    CU(cudaMemcpyAsync(mine.data(), h->list_size.p, nc * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(all.data(), list_size_dev, nc * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    uint64_t total = 0;
    for (size_t c = 0; c < nc; c++) {
        total += all[c];
        if (c >= h->res_begin && c < h->res_end && all[c] != mine[c])
            return h->fail(IVFHNSW_B200_EINVAL, "set_list_meta: resident list %zu has %u records here, %u in the new table", c,
                           mine[c], all[c]);
    }
    CU(cudaMemcpyAsync(h->list_size.p, list_size_dev, nc * 4, cudaMemcpyDeviceToDevice, st));
    if (nsubc) {
        CU(cudaMemcpyAsync(h->sub_sizes.p, sub_sizes_dev, nc * nsubc * 4, cudaMemcpyDeviceToDevice, st));
        CU(cudaMemcpyAsync(h->alphas.p, alphas_dev, nc * 4, cudaMemcpyDeviceToDevice, st));
    }
    CU(cudaStreamSynchronize(st));
    h->ntotal = (size_t)total;
    return 0;
}

int ivfhnsw_b200_ntotal(ivfhnsw_b200_t *h, size_t *ntotal)
{
    *ntotal = h->st_dirty ? h->st_n + h->rs_n : h->ntotal;
    return 0;
}

int ivfhnsw_b200_export_lists(ivfhnsw_b200_t *h, int64_t *list_off, uint32_t *ids, uint8_t *codes, uint8_t *norm_codes,
                              uint32_t *nn_idx, uint32_t *sub_sizes, float *alphas, float *icd)
{
    if (int r = ivfhnsw_b200_finalize(h)) return r;
    std::lock_guard<std::mutex> lk(h->mu);
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    if (!h->have_lists) return h->fail(IVFHNSW_B200_ESTATE, "no lists");
    if (h->res_begin != 0 || h->res_end != h->nc)
        return h->fail(IVFHNSW_B200_ESTATE, "cannot export: only the lists [%u, %u) are resident on this handle", h->res_begin,
                       h->res_end);
    const size_t nc = h->nc, n = h->ntotal, M = h->M, nsubc = h->nsubc;
    cudaStream_t st = h->stream;
    std::vector<uint32_t> ro(nc + 1);
    CU(cudaMemcpyAsync(ro.data(), h->raw_off.p, (nc + 1) * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    for (size_t i = 0; i <= nc; i++) list_off[i] = ro[i];
    if (n) {
        CU(h->w_a.ensure(n * M));
        CU(h->w_b.ensure(n));
        CU(h->w_c.ensure(n * 4));
        CU(launch_gather_lists(st, LAUNCHES, (int)nc, (int)M, h->list_size.as<uint32_t>(), h->list_slot.as<uint32_t>(),
                               h->raw_off.as<uint32_t>(), h->codes.as<uint8_t>(), h->norm_codes.as<uint8_t>(),
                               h->ids.as<uint32_t>(), h->w_a.as<uint8_t>(), h->w_b.as<uint8_t>(), h->w_c.as<uint32_t>()));
        CU(cudaMemcpyAsync(codes, h->w_a.p, n * M, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(norm_codes, h->w_b.p, n, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(ids, h->w_c.p, n * 4, cudaMemcpyDeviceToHost, st));
    }
    if (nsubc) {
        if (nn_idx) CU(cudaMemcpyAsync(nn_idx, h->nn_idx.p, nc * nsubc * 4, cudaMemcpyDeviceToHost, st));
        if (sub_sizes) CU(cudaMemcpyAsync(sub_sizes, h->sub_sizes.p, nc * nsubc * 4, cudaMemcpyDeviceToHost, st));
        if (alphas) CU(cudaMemcpyAsync(alphas, h->alphas.p, nc * 4, cudaMemcpyDeviceToHost, st));
        if (icd) CU(cudaMemcpyAsync(icd, h->icd.p, nc * nsubc * 4, cudaMemcpyDeviceToHost, st));
    }
    CU(cudaStreamSynchronize(st));
    return 0;
}

// ------------------------------------------------------------------------------------ search
static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

static DeviceLists device_lists(ivfhnsw_b200_t *h)
{
    const bool grouping = h->nsubc > 0;
    DeviceLists L;
    L.list_size = h->list_size.as<uint32_t>(); L.list_slot = h->list_slot.as<uint32_t>();
    L.codes = h->codes.as<uint8_t>(); L.norm_codes = h->norm_codes.as<uint8_t>();
    L.sub_ids = grouping ? h->sub_ids.as<uint8_t>() : nullptr; L.ids = h->ids.as<uint32_t>();
    L.nn_idx = h->nn_idx.as<uint32_t>(); L.sub_sizes = h->sub_sizes.as<uint32_t>();
    L.alphas = h->alphas.as<float>(); L.icd = h->icd.as<float>();
    L.centroid_norms = h->centroid_norms.as<float>(); L.centroids = h->centroids.as<float>();
    return L;
}

static int search_checks(ivfhnsw_b200_t *h, size_t k)
{
    if (!h->have_lists || h->st_dirty) return h->fail(IVFHNSW_B200_ESTATE, "lists not built: call finalize()/read() first");
    if (!h->have_pq || !h->have_norm) return h->fail(IVFHNSW_B200_ESTATE, "PQ codebooks not set");
    if (!h->have_centroids) return h->fail(IVFHNSW_B200_ESTATE, "centroids not set");
    if (h->do_opq && !h->have_opq) return h->fail(IVFHNSW_B200_ESTATE, "do_opq set but no OPQ matrix");
    if (k == 0 || k > 2048) return h->fail(IVFHNSW_B200_EINVAL, "k %zu out of range [1, 2048]", k);
    if (h->nsubc && h->nprobe * h->nsubc + h->nprobe + 1 > 65535)
        return h->fail(IVFHNSW_B200_EINVAL, "Grouping search: nprobe * (nsubc + 1) = %zu exceeds 65534 (the pruning pass indexes "
                       "its (probe, sub-group) rows with 16 bits); lower nprobe", h->nprobe * (h->nsubc + 1));
    const size_t smem = scan_smem_bytes((int)h->M, (int)h->nprobe, (int)k, (int)h->nsubc);
    if (smem > 227 * 1024) return h->fail(IVFHNSW_B200_EINVAL, "nprobe/k too large for the scan kernel (%zu B smem)", smem);
    return 0;
}

static size_t query_batch(const ivfhnsw_b200_t *h, size_t nq, size_t k)
{
    const size_t rec = seg_record_words((int)h->nprobe, (int)h->nsubc) * 4;
    const size_t per_q = 2 * rec + 32 * k * 12 + h->d * 8 + h->nprobe * 8 + (size_t)1024 * h->M +
                         (h->nsubc ? g1_work_bytes((int)h->nprobe, (int)h->nsubc) : 0);
    size_t qb = h->query_batch > 0 ? (size_t)h->query_batch : 16384;
    qb = std::min(qb, std::max<size_t>(64, ((size_t)3 << 30) / per_q));
    return std::max<size_t>(1, std::min(qb, nq));
}

// Stage 1 of a search: OPQ rotation, coarse probes (unless given) and the work tables of queries
// [q_begin, q_end) of the batch x_dev[nq][d], written as records q_begin.. of `tables` (kernels.h SegTables).
// The records do not depend on the shard.
static int prepare_core(ivfhnsw_b200_t *h, size_t q_begin, size_t q_end, const float *x_dev, const float *cd_dev,
                        const uint32_t *ci_dev, uint32_t *tables)
{
    const int d = (int)h->d, nprobe = (int)h->nprobe, nsubc = (int)h->nsubc;
    const bool grouping = nsubc > 0;
    cudaStream_t st = h->stream;
    if (q_end <= q_begin) return 0;
    const size_t nq = q_end - q_begin;
    const size_t qb = query_batch(h, nq, 10);
    const size_t stride = seg_record_words(nprobe, nsubc);
    CU(h->w_seg.ensure(qb * 12 + 1024));
    if (grouping) CU(h->w_qsd.ensure(qb * (size_t)nprobe * nsubc * 4 + 256));
    if (grouping) CU(h->w_g1.ensure(qb * g1_work_bytes(nprobe, nsubc) + 256));
    if (!cd_dev) {
        CU(h->w_cd.ensure(qb * nprobe * 4));
        CU(h->w_ci.ensure(qb * nprobe * 4));
    }
    if (h->do_opq) CU(h->w_xrot.ensure(qb * d * 4));
    const DeviceLists L = device_lists(h);
    h->last_thr.clear();
    h->last_scanned.clear();
    for (size_t q0 = q_begin; q0 < q_end; q0 += qb) {
        const int nb = (int)std::min(qb, q_end - q0);
        SegTables T;
        T.rec = tables + q0 * stride;
        T.stride = stride; T.nprobe = nprobe; T.nsubc = nsubc;
        T.qsd = grouping ? h->w_qsd.as<float>() : nullptr;
        T.threshold = h->w_seg.as<float>();
        T.nscanned = reinterpret_cast<uint32_t *>(T.threshold + qb);
        T.nwalked = T.nscanned + qb;
        const float *xq = x_dev + q0 * d;
        if (h->timing) CU(cudaEventRecord(h->ev[0], st));
        if (h->do_opq) {   // IndexIVF_HNSW.cpp:240
            CU(launch_rotate(st, LAUNCHES, nb, d, h->opqA.as<float>(), false, xq, h->w_xrot.as<float>()));
            xq = h->w_xrot.as<float>();
        }
        const float *cd;
        const uint32_t *ci;
        if (cd_dev) { cd = cd_dev + q0 * nprobe; ci = ci_dev + q0 * nprobe; }
        else {
            if (int r = ivfb200_coarse_dev(h, nb, xq, nprobe, 0, h->w_cd.as<float>(), h->w_ci.as<uint32_t>())) return r;
            cd = h->w_cd.as<float>(); ci = h->w_ci.as<uint32_t>();
        }
        if (h->timing) CU(cudaEventRecord(h->ev[1], st));
        if (grouping)
            CU(launch_prep_grouping(st, LAUNCHES, nb, d, nprobe, nsubc, h->max_codes, h->do_pruning, xq, cd, ci, L, T,
                                    h->w_g1.p,
                                    h->timing ? h->ev[8] : nullptr, h->timing ? h->ev[4] : nullptr,
                                    h->counters.as<unsigned long long>() + 5));
        else
            CU(launch_prep_adc(st, LAUNCHES, nb, nprobe, h->max_codes, cd, ci, L, T));
        reduce_u32_kernel<<<grid_for(nb, 256, 64), 256, 0, st>>>(nb, T.nscanned, T.nwalked,
                                                                h->counters.as<unsigned long long>());
        ++h->stats.kernel_launches;
        if (h->timing) {
            CU(cudaEventRecord(h->ev[2], st));
            CU(cudaStreamSynchronize(st));
            float ms = 0;
            CU(cudaEventElapsedTime(&ms, h->ev[0], h->ev[1])); h->stats.coarse_ms += ms;
            CU(cudaEventElapsedTime(&ms, h->ev[1], h->ev[2])); h->stats.prep_ms += ms;
            if (grouping) { CU(cudaEventElapsedTime(&ms, h->ev[8], h->ev[4])); h->stats.g1_ms += ms; }
            h->stats.prep_launches += 1;
        }
        if (q0 == q_begin && h->debug) {   // keep debug outputs of the first batch (copied lazily by last_thresholds)
            h->last_thr.resize(nb);
            h->last_scanned.resize(nb);
            CU(cudaMemcpyAsync(h->last_thr.data(), T.threshold, (size_t)nb * 4, cudaMemcpyDeviceToHost, st));
            CU(cudaMemcpyAsync(h->last_scanned.data(), T.nscanned, (size_t)nb * 4, cudaMemcpyDeviceToHost, st));
            if (nq > qb) CU(cudaStreamSynchronize(st));  // the scratch is reused by the next batch
        }
    }
    return 0;
}

// Stage 2: lookup tables + ADC scan of THIS GPU's lists for all nq queries, driven by the records.
static int scan_core(ivfhnsw_b200_t *h, size_t nq, size_t k, const float *x_dev, const uint32_t *tables, float *out_d,
                     int64_t *out_l, uint32_t *out_seq, bool xrot_ready = false, const uint32_t *qoff = nullptr,
                     const uint32_t *qnp = nullptr)
{
    if (nq == 0) return 0;
    const int d = (int)h->d, M = (int)h->M, nprobe = (int)h->nprobe, nsubc = (int)h->nsubc;
    const bool grouping = nsubc > 0;
    cudaStream_t st = h->stream;
    const size_t qb = query_batch(h, nq, k);
    const size_t stride = seg_record_words(nprobe, nsubc);
    int nsplit = (int)h->scan_split;
    if (nsplit <= 0) {
        const size_t target = 148 * 6;
        nsplit = (int)std::min<size_t>(32, (target + qb - 1) / qb);
    }
    nsplit = std::max(1, std::min(nsplit, 32));
    if (nsplit > 1) CU(h->w_part.ensure(qb * nsplit * k * 12 + 256));
    CU(h->w_lut.ensure(qb * 256 * M * 4 + 256));
    if (h->do_opq) CU(h->w_xrot.ensure(qb * d * 4));
    const DeviceLists L = device_lists(h);
    for (size_t q0 = 0; q0 < nq; q0 += qb) {
        const int nb = (int)std::min(qb, nq - q0);
        SegTables T;
        T.rec = const_cast<uint32_t *>(tables) + (qoff ? 0 : q0 * stride);
        T.stride = stride; T.nprobe = nprobe; T.nsubc = nsubc;
        T.qsd = nullptr; T.threshold = nullptr; T.nscanned = nullptr; T.nwalked = nullptr;
        T.qoff = qoff ? qoff + q0 : nullptr;
        T.qnp = qnp ? qnp + q0 : nullptr;
        const float *xq = x_dev + q0 * d;
        if (h->timing) CU(cudaEventRecord(h->ev[2], st));
        if (h->do_opq) {
            // xrot_ready: prepare_core just rotated these very queries into w_xrot (one batch)
            if (!(xrot_ready && nq <= qb))
                CU(launch_rotate(st, LAUNCHES, nb, d, h->opqA.as<float>(), false, xq, h->w_xrot.as<float>()));
            xq = h->w_xrot.as<float>();
        }
        float *pd = nullptr;
        uint32_t *ps = nullptr, *pl = nullptr;
        if (nsplit > 1) {
            pd = h->w_part.as<float>();
            ps = reinterpret_cast<uint32_t *>(pd + (size_t)qb * nsplit * k);
            pl = ps + (size_t)qb * nsplit * k;
        }
        CU(launch_lut(st, LAUNCHES, M, nb, d, xq, h->pqc.as<float>(), h->w_lut.as<float>(), T.qnp));
        if (h->debug >= 2) CU(cudaStreamSynchronize(st));   // localise asynchronous kernel faults
        if (h->timing) CU(cudaEventRecord(h->ev[5], st));
        CU(launch_scan(st, LAUNCHES, M, grouping, nb, d, nprobe, nsubc, h->do_pruning, (int)k, nsplit, h->w_lut.as<float>(),
                       h->norm_table.as<float>(), L, T, (uint32_t)h->c_begin, (uint32_t)h->c_end, pd, ps, pl,
                       out_d + q0 * k, out_l + q0 * k, out_seq ? out_seq + q0 * k : nullptr));
        if (h->debug >= 2) CU(cudaStreamSynchronize(st));
        if (h->timing) CU(cudaEventRecord(h->ev[3], st));
        if (nsplit > 1)
            CU(launch_merge(st, LAUNCHES, nb, (int)k, nsplit, pd, ps, pl, L.ids, out_d + q0 * k, out_l + q0 * k,
                            out_seq ? out_seq + q0 * k : nullptr));
        if (h->timing) {
            CU(cudaStreamSynchronize(st));
            float ms = 0;
            CU(cudaEventElapsedTime(&ms, h->ev[2], h->ev[3])); h->stats.scan_ms += ms;
            CU(cudaEventElapsedTime(&ms, h->ev[2], h->ev[5])); h->stats.lut_ms += ms;
            h->stats.scan_launches += 1;
        }
    }
    h->stats.queries += nq;
    return 0;
}

static int search_core(ivfhnsw_b200_t *h, size_t nq, size_t k, const float *x_dev, const float *cd_dev,
                       const uint32_t *ci_dev, float *out_d, int64_t *out_l, uint32_t *out_seq = nullptr)
{
    if (int r = search_checks(h, k)) return r;
    if (nq == 0) return 0;
    // batch-wise so that the internal record buffer stays bounded
    const size_t qb = query_batch(h, nq, k);
    const size_t stride = seg_record_words((int)h->nprobe, (int)h->nsubc);
    CU(h->w_sub_base.ensure(qb * stride * 4 + 256));
    uint32_t *tables = h->w_sub_base.as<uint32_t>();
    for (size_t q0 = 0; q0 < nq; q0 += qb) {
        const size_t nb = std::min(qb, nq - q0);
        if (int r = prepare_core(h, 0, nb, x_dev + q0 * h->d, cd_dev ? cd_dev + q0 * h->nprobe : nullptr,
                                 ci_dev ? ci_dev + q0 * h->nprobe : nullptr, tables))
            return r;
        const bool xrot_ready = h->do_opq && nb <= query_batch(h, nb, 10);   // prepare_core ran ONE batch
        if (int r = scan_core(h, nb, k, x_dev + q0 * h->d, tables, out_d + q0 * k, out_l + q0 * k,
                              out_seq ? out_seq + q0 * k : nullptr, xrot_ready))
            return r;
    }
    return 0;
}

int ivfhnsw_b200_search_given_coarse_dev(ivfhnsw_b200_t *h, size_t nq, size_t k, const float *x_dev,
                                         const float *cd_dev, const uint32_t *ci_dev, float *dd, int64_t *ld)
{
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    return search_core(h, nq, k, x_dev, cd_dev, ci_dev, dd, ld);
}
int ivfhnsw_b200_search_dev(ivfhnsw_b200_t *h, size_t nq, size_t k, const float *x_dev, float *dd, int64_t *ld)
{
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    return search_core(h, nq, k, x_dev, nullptr, nullptr, dd, ld);
}

int ivfhnsw_b200_search_partial_dev(ivfhnsw_b200_t *h, size_t nq, size_t k, const float *x_dev, const float *cd_dev,
                                    const uint32_t *ci_dev, float *part_d, uint32_t *part_seq, int64_t *part_labels)
{
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    return search_core(h, nq, k, x_dev, cd_dev, ci_dev, part_d, part_labels, part_seq);
}
int ivfhnsw_b200_tables_bytes(ivfhnsw_b200_t *h, size_t *bytes_per_query)
{
    *bytes_per_query = seg_record_words((int)h->nprobe, (int)h->nsubc) * 4;
    return 0;
}
int ivfhnsw_b200_prepare_dev(ivfhnsw_b200_t *h, size_t nq, const float *x_dev, size_t q_begin, size_t q_end,
                             void *tables_dev)
{
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    if (int r = search_checks(h, 1)) return r;
    if (q_begin > q_end || q_end > nq) return h->fail(IVFHNSW_B200_EINVAL, "query slice [%zu, %zu) outside [0, %zu)", q_begin, q_end, nq);
    return prepare_core(h, q_begin, q_end, x_dev, nullptr, nullptr, reinterpret_cast<uint32_t *>(tables_dev));
}
int ivfhnsw_b200_scan_tables_dev(ivfhnsw_b200_t *h, size_t nq, size_t k, const float *x_dev, const void *tables_dev,
                                 float *part_d, uint32_t *part_seq, int64_t *part_labels)
{
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    if (int r = search_checks(h, k)) return r;
    return scan_core(h, nq, k, x_dev, reinterpret_cast<const uint32_t *>(tables_dev), part_d, part_labels, part_seq);
}
int ivfhnsw_b200_pack_records_dev(ivfhnsw_b200_t *h, size_t nq, const void *tables_dev, size_t nparts, void *packed_dev,
                                  uint32_t *cnt_dev, uint32_t *off_dev, uint64_t *part_words_dev)
{
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    if (nparts == 0 || nparts > 64) return h->fail(IVFHNSW_B200_EINVAL, "nparts out of range [1, 64]");
    const size_t stride = seg_record_words((int)h->nprobe, (int)h->nsubc);
    if (nq * stride >= (1ull << 32)) return h->fail(IVFHNSW_B200_EINVAL, "pack_records: batch too large (32-bit word offsets)");
    static_assert(sizeof(unsigned long long) == sizeof(uint64_t), "64-bit counters");
    CU(launch_pack_records(h->stream, LAUNCHES, (int)nq, (int)h->nprobe, (int)h->nsubc, (int)h->nc, (int)nparts,
                           reinterpret_cast<const uint32_t *>(tables_dev), stride, reinterpret_cast<uint32_t *>(packed_dev),
                           cnt_dev, off_dev, reinterpret_cast<unsigned long long *>(part_words_dev)));
    return 0;
}
int ivfhnsw_b200_scan_packed_dev(ivfhnsw_b200_t *h, size_t nq, size_t k, const float *x_dev, const void *packed_dev,
                                 const uint32_t *qoff_dev, const uint32_t *qnp_dev, float *part_d, uint32_t *part_seq,
                                 int64_t *part_labels)
{
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    if (int r = search_checks(h, k)) return r;
    if (!qoff_dev || !qnp_dev) return h->fail(IVFHNSW_B200_EINVAL, "scan_packed: per-query offsets / probe counts missing");
    return scan_core(h, nq, k, x_dev, reinterpret_cast<const uint32_t *>(packed_dev), part_d, part_labels, part_seq, false,
                     qoff_dev, qnp_dev);
}
int ivfhnsw_b200_merge_partials_dev(ivfhnsw_b200_t *h, size_t nq, size_t k, size_t nparts, const float *part_d,
                                    const uint32_t *part_seq, const int64_t *part_labels, float *out_d,
                                    int64_t *out_labels)
{
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    if (nparts == 0 || nparts > 64) return h->fail(IVFHNSW_B200_EINVAL, "nparts out of range");
    CU(launch_merge_shards(h->stream, LAUNCHES, (int)nq, (int)k, (int)nparts, part_d, part_seq, part_labels, out_d,
                           out_labels));
    return 0;
}

static int search_host(ivfhnsw_b200_t *h, size_t nq, size_t k, const float *x, const float *cd, const uint32_t *ci,
                       float *distances, int64_t *labels)
{
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    if (nq == 0) return 0;
    cudaStream_t st = h->stream;
    const size_t np = h->nprobe;
    CU(h->w_x.ensure(nq * h->d * 4));
    CU(h->w_out_d.ensure(nq * k * 4));
    CU(h->w_out_l.ensure(nq * k * 8));
    CU(cudaMemcpyAsync(h->w_x.p, x, nq * h->d * 4, cudaMemcpyHostToDevice, st));
    const float *cdd = nullptr;
    const uint32_t *cid = nullptr;
    if (cd) {
        CU(h->w_e.ensure(nq * np * 4));
        CU(h->w_f.ensure(nq * np * 4));
        CU(cudaMemcpyAsync(h->w_e.p, cd, nq * np * 4, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(h->w_f.p, ci, nq * np * 4, cudaMemcpyHostToDevice, st));
        cdd = h->w_e.as<float>();
        cid = h->w_f.as<uint32_t>();
    }
    if (int r = search_core(h, nq, k, h->w_x.as<float>(), cdd, cid, h->w_out_d.as<float>(), h->w_out_l.as<int64_t>()))
        return r;
    CU(cudaMemcpyAsync(distances, h->w_out_d.p, nq * k * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(labels, h->w_out_l.p, nq * k * 8, cudaMemcpyDeviceToHost, st));
    if (cd) { CU(cudaStreamSynchronize(st)); return 0; }
    return check_coarse_flags(h);
}

int ivfhnsw_b200_search(ivfhnsw_b200_t *h, size_t nq, size_t k, const float *x, float *distances, int64_t *labels)
{
    return search_host(h, nq, k, x, nullptr, nullptr, distances, labels);
}
int ivfhnsw_b200_search_given_coarse(ivfhnsw_b200_t *h, size_t nq, size_t k, const float *x, const float *cd,
                                     const uint32_t *ci, float *distances, int64_t *labels)
{
    if (!cd || !ci) return h->fail(IVFHNSW_B200_EINVAL, "coarse candidates missing");
    return search_host(h, nq, k, x, cd, ci, distances, labels);
}

int ivfhnsw_b200_rerank(ivfhnsw_b200_t *h, size_t nq, size_t ncand, size_t k, const float *queries, const float *cand,
                        const int64_t *cand_labels, float *distances, int64_t *labels)
{
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    if (nq == 0) return 0;
    if (k == 0 || k > ncand || ncand > 8192) return h->fail(IVFHNSW_B200_EINVAL, "rerank: need 1 <= k <= ncand <= 8192");
    cudaStream_t st = h->stream;
    const size_t d = h->d;
    CU(h->w_x.ensure(nq * d * 4));
    CU(h->w_a.ensure(nq * ncand * d * 4));
    CU(h->w_e.ensure(nq * ncand * 8));
    CU(h->w_out_d.ensure(nq * k * 4));
    CU(h->w_out_l.ensure(nq * k * 8));
    CU(cudaMemcpyAsync(h->w_x.p, queries, nq * d * 4, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(h->w_a.p, cand, nq * ncand * d * 4, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(h->w_e.p, cand_labels, nq * ncand * 8, cudaMemcpyHostToDevice, st));
    CU(launch_rerank(st, LAUNCHES, (int)nq, (int)ncand, (int)k, (int)d, h->w_x.as<float>(), h->w_a.as<float>(),
                     h->w_e.as<int64_t>(), h->w_out_d.as<float>(), h->w_out_l.as<int64_t>()));
    CU(cudaMemcpyAsync(distances, h->w_out_d.p, nq * k * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(labels, h->w_out_l.p, nq * k * 8, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    return 0;
}

int ivfhnsw_b200_get_stats(ivfhnsw_b200_t *h, ivfhnsw_b200_stats_t *out)
{
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    unsigned long long c[8] = {0};
    CU(cudaStreamSynchronize(h->stream));
    CU(cudaMemcpy(c, h->counters.p, 64, cudaMemcpyDeviceToHost));
    h->stats.scanned_codes = c[0];
    h->stats.walked_codes = c[1];
    h->stats.g1_rows = c[5];
    *out = h->stats;
    if (c[4] & 0xffffffffull)
        return h->fail(IVFHNSW_B200_ECUDA, "coarse select overflow flags 0x%llx (raise coarse_slack)", c[4] & 0xffffffffull);
    return 0;
}
int ivfhnsw_b200_reset_stats(ivfhnsw_b200_t *h)
{
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    CU(cudaStreamSynchronize(h->stream));
    CU(cudaMemset(h->counters.p, 0, 64));
    h->stats = ivfhnsw_b200_stats_t{};
    return 0;
}
int ivfhnsw_b200_last_thresholds(ivfhnsw_b200_t *h, size_t cap, float *thresholds, uint32_t *scanned)
{
    if (bind(h)) return IVFHNSW_B200_ECUDA;
    CU(cudaStreamSynchronize(h->stream));
    const size_t n = std::min(cap, h->last_thr.size());
    if (thresholds) memcpy(thresholds, h->last_thr.data(), n * 4);
    if (scanned) memcpy(scanned, h->last_scanned.data(), n * 4);
    return (int)n;
}

}  // extern "C"
