// <ivf-hnsw/orcv.h> of the B200 drop-in: the fixed header record of the ORCV directory layout that
// IndexIVF_HNSW::write2 produces (reference orcv.h:11-22, written to <home_dir>/hdr.vec).  Same field names, order and
// sizes as the reference's record (the fork's mains fill and read it member by member).
#pragma once
#include <stdint.h>

struct orcvhdr {
    uint32_t n, nc;                   // vectors in the index, centroids
    uint32_t code_size, code_bytes;   // PQ code size as passed to the constructor; code_size / 8 as the reference computes it
    uint32_t d;                       // vector dimension
    uint32_t M, efConstruction;       // HNSW parameters (carried, unused here: the coarse stage is an exact GEMM)
    float dmatch, dnear;              // "match" / "near match" distance thresholds
    uint8_t do_opq;                   // OPQ rotation in use
};
typedef struct orcvhdr orcvhdr_t;
static_assert(sizeof(orcvhdr_t) == 40, "ORCV header record is 9 words + 1 byte, padded to 40");
