// <ivf-hnsw/orcv.h> of the B200 drop-in: the fixed header record of the ORCV directory layout that
// IndexIVF_HNSW::write2 produces (reference orcv.h:11-22, written to <home_dir>/hdr.vec).
#pragma once
#include <stdint.h>

typedef struct orcvhdr {
    uint32_t n;               // vectors in the index
    uint32_t nc;              // centroids
    uint32_t code_size;       // PQ code size as passed to the constructor
    uint32_t code_bytes;      // code_size / 8, as the reference computes it
    uint32_t d;               // vector dimension
    uint32_t M;               // HNSW: edges per point
    uint32_t efConstruction;  // HNSW: construction beam
    float dmatch;             // "match" distance threshold
    float dnear;              // "near match" distance threshold
    uint8_t do_opq;           // OPQ rotation in use
} orcvhdr_t;
