// Minimal faiss::ProductQuantizer for the ivf-hnsw drop-in: the public data members the reference's
// callers read or replace (faiss/ProductQuantizer.h:24-60) and nothing of the CPU arithmetic -- encoding,
// decoding and table building run on the GPU behind the C ABI.
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>

namespace faiss {

struct ProductQuantizer {
    size_t d;               // size of the input vectors
    size_t M;               // number of sub-quantizers
    size_t nbits;           // bits per quantization index
    size_t dsub;            // dimensionality of each sub-vector
    size_t byte_per_idx;    // bytes per index component (1 for nbits = 8)
    size_t code_size;       // bytes per indexed vector
    size_t ksub;            // centroids per sub-quantizer
    bool verbose;
    std::vector<float> centroids;   // (M, ksub, dsub)

    ProductQuantizer(size_t d_ = 0, size_t M_ = 1, size_t nbits_ = 8) : d(d_), M(M_), nbits(nbits_), verbose(false)
    {
        set_derived_values();
    }
    void set_derived_values()
    {
        dsub = M ? d / M : 0;
        byte_per_idx = (nbits + 7) / 8;
        code_size = byte_per_idx * M;
        ksub = (size_t)1 << nbits;
        centroids.resize(d * ksub);
    }
    float *get_centroids(size_t m, size_t i) { return &centroids[(m * ksub + i) * dsub]; }
    const float *get_centroids(size_t m, size_t i) const { return &centroids[(m * ksub + i) * dsub]; }
};

}  // namespace faiss
