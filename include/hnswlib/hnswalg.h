// Drop-in stand-in for the reference's <hnswlib/hnswalg.h> (reference hnswlib/hnswalg.h:45,60-170).
//
// In the B200 build the coarse quantizer is the EXACT tensor-core GEMM search over the centroid table
// (ivfhnsw_b200_coarse), so there is no graph to traverse.  The type only keeps the members the
// reference's tests touch (`index->quantizer->efSearch = 220;`, tests/test_ivfhnsw_sift1b.cpp:108,183) so
// that they compile unchanged; efSearch has no effect on an exact search.
#pragma once
#include <cstddef>
#include <cstdint>

namespace hnswlib {

typedef unsigned int idx_t;   // the same type as ivfhnsw::IndexIVF_HNSW::idx_t

struct HierarchicalNSW {
    size_t d_;            // vector dimension
    size_t maxelements_;  // number of centroids
    size_t M_;
    size_t efConstruction_;
    size_t efSearch;      // accepted and ignored: the search is exhaustive and exact

    HierarchicalNSW(size_t d, size_t maxelements, size_t M = 16, size_t efConstruction = 500)
        : d_(d), maxelements_(maxelements), M_(M), efConstruction_(efConstruction), efSearch(0) {}
};

}  // namespace hnswlib
