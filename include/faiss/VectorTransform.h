// Minimal faiss::VectorTransform / LinearTransform for the ivf-hnsw drop-in (faiss/VectorTransform.h:26-110):
// the OPQ rotation is a plain d x d matrix; apply() exists for callers that rotate on the host, the
// index itself rotates queries and centroids on the GPU.
#pragma once
#include <cstddef>
#include <vector>

namespace faiss {

struct VectorTransform {
    int d_in, d_out;
    bool is_trained;
    explicit VectorTransform(int d_in_ = 0, int d_out_ = 0) : d_in(d_in_), d_out(d_out_), is_trained(true) {}
    virtual ~VectorTransform() {}
    // returns a new float[n * d_out] the caller deletes (reference IndexIVF_HNSW.cpp:240,295)
    float *apply(long n, const float *x) const
    {
        float *xt = new float[(size_t)n * d_out];
        apply_noalloc(n, x, xt);
        return xt;
    }
    virtual void apply_noalloc(long n, const float *x, float *xt) const = 0;
};

struct LinearTransform : VectorTransform {
    bool have_bias;
    std::vector<float> A;   // row-major d_out x d_in
    std::vector<float> b;
    bool verbose;
    explicit LinearTransform(int d_in_ = 0, int d_out_ = 0, bool have_bias_ = false)
        : VectorTransform(d_in_, d_out_), have_bias(have_bias_), verbose(false) {}
    void apply_noalloc(long n, const float *x, float *xt) const override;
    void transform_transpose(long n, const float *y, float *x) const;
};

}  // namespace faiss
