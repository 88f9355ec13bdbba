// Handle state shared by api.cu and train.cu.
#pragma once
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/ivfhnsw_b200.h"
#include "kernels.h"

namespace ivfb200 {
struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    DevBuf() = default;
    DevBuf(const DevBuf &) = delete;              // owns its allocation: locals release it on every exit path
    DevBuf &operator=(const DevBuf &) = delete;
    ~DevBuf() { release(); }
    cudaError_t ensure(size_t bytes)
    {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e == cudaSuccess) cap = bytes;
        return e;
    }
    // keep the first `keep` bytes
    cudaError_t grow(size_t bytes, size_t keep, cudaStream_t st)
    {
        if (bytes <= cap) return cudaSuccess;
        size_t ncap = std::max(bytes, cap + cap / 2);
        void *np = nullptr;
        cudaError_t e = cudaMalloc(&np, ncap);
        if (e != cudaSuccess) return e;
        if (p && keep) {
            e = cudaMemcpyAsync(np, p, keep, cudaMemcpyDeviceToDevice, st);
            if (e != cudaSuccess) return e;
            e = cudaStreamSynchronize(st);
            if (e != cudaSuccess) return e;
        }
        if (p) cudaFree(p);
        p = np;
        cap = ncap;
        return cudaSuccess;
    }
    void release()
    {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    template <typename T> T *as() const { return reinterpret_cast<T *>(p); }
};

}  // namespace ivfb200

using ivfb200::DevBuf;

struct ivfhnsw_b200 {
    int device = 0;
    size_t d = 0, nc = 0, M = 0, nsubc = 0;
    size_t nprobe = 1, max_codes = 1000000000;
    bool do_opq = false, do_pruning = false;
    uint32_t c_begin = 0, c_end = 0;       // lists this handle scans
    uint32_t res_begin = 0, res_end = 0;   // lists whose codes are resident (the shard at CSR-build time)
    cudaStream_t stream = nullptr;
    bool own_stream = true;
    cudaEvent_t ev[10] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    std::string err;
    std::mutex mu;

    // model
    DevBuf centroids, centroid_norms, cnorm_gemm, cent_f16, pqc, norm_table, opqA;
    bool have_centroids = false, have_pq = false, have_norm = false, have_opq = false, have_cnorms = false;
    bool gemm_norms_valid = false, f16_valid = false;
    // lists (device CSR in slot space)
    DevBuf list_size, list_slot, raw_off, codes, norm_codes, sub_ids, ids;
    DevBuf nn_idx, sub_sizes, alphas, icd;
    bool have_lists = false, have_group_meta = false;
    size_t ntotal = 0, nslots = 0;
    // staging (records not yet in the CSR)
    DevBuf st_list, st_ids, st_codes, st_norm, st_sub;
    size_t st_n = 0;          // encoded records staged
    // Grouping: raw points whose group is not encoded yet (list, id, vector); flushed by flush_groups / finalize
    DevBuf rs_list, rs_ids, st_x, grp_seen;
    size_t rs_n = 0;
    bool grp_seen_valid = false;
    bool st_dirty = false;
    // workspace
    DevBuf w_x, w_xrot, w_cd, w_ci, w_seg, w_sub_base, w_qsd, w_g1, w_part, w_lut, w_coarse, w_xf16, w_out_d, w_out_l;
    DevBuf w_a, w_b, w_c, w_d2, w_e, w_f, w_cub, w_small;
    DevBuf counters;  // [0] scanned, [1] walked
    // options / stats
    long coarse_slack = 8, coarse_impl = 3, scan_split = 0, query_batch = 0, debug = 0;
    long long stage_bytes = 4ll << 30;   // raw Grouping points staged before add_group encodes them (option "stage_bytes")
    bool timing = false;
    ivfhnsw_b200_stats_t stats{};
    std::vector<float> last_thr;
    std::vector<uint32_t> last_scanned;

    int fail(int code, const char *fmt, ...)
    {
        char buf[512];
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(buf, sizeof buf, fmt, ap);
        va_end(ap);
        err = buf;
        return code;
    }
};

#define CU(call)                                                                                         \
    do {                                                                                                 \
        cudaError_t e__ = (call);                                                                        \
        if (e__ != cudaSuccess)                                                                          \
            return h->fail(IVFHNSW_B200_ECUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
    } while (0)

#define LAUNCHES (&h->stats.kernel_launches)


extern "C" {
#define IVF_HIDDEN __attribute__((visibility("hidden")))
// shared host-side stages (api.cu), internal linkage-by-visibility
IVF_HIDDEN int ivfb200_coarse_dev(ivfhnsw_b200_t *h, size_t nq, const float *xq, size_t nprobe, int exclude_self, float *out_d,
               uint32_t *out_i);
IVF_HIDDEN int ivfb200_assign_dev(ivfhnsw_b200_t *h, size_t n, const float *x_dev, uint32_t *labels_dev);
IVF_HIDDEN int ivfb200_ensure_group_meta(ivfhnsw_b200_t *h);
// group stage of add_group / train_pq: alphas_out[nc] and sub_out[n] for n device points with list ids
IVF_HIDDEN int ivfb200_group_stage(ivfhnsw_b200_t *h, size_t n, const float *x_dev, const uint32_t *list_dev, float *alphas_out,
                uint8_t *sub_out, int zero_alphas, uint32_t *seen);
IVF_HIDDEN int ivfb200_encode_dev(ivfhnsw_b200_t *h, size_t n, const float *x_dev, const uint32_t *idx_dev, const uint8_t *sub_dev,
               const float *alphas_dev, uint8_t *codes_dev, uint8_t *norm_dev, float *norms_out, float *residual_out);
}  // extern "C"
