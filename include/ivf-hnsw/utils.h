// <ivf-hnsw/utils.h> of the B200 drop-in: the host-side helpers the reference's mains use
// (reference utils.h:27-127): timers, .bvecs/.fvecs/.ivecs readers, random_subset.
// Written from scratch; semantics (names, argument order, "file error" + exit(1) on a dimension
// mismatch) follow the reference so that tests/*.cpp compile and behave the same.
#pragma once
#include <sys/stat.h>

#include <cassert>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

#define EPS 0.00001

namespace ivfhnsw {

// wall-clock stopwatch, microsecond resolution (reference utils.h:27-39)
class StopW {
    std::chrono::steady_clock::time_point start_;

public:
    StopW() { reset(); }
    float getElapsedTimeMicro()
    {
        const auto now = std::chrono::steady_clock::now();
        return (float)std::chrono::duration_cast<std::chrono::microseconds>(now - start_).count();
    }
    void reset() { start_ = std::chrono::steady_clock::now(); }
};

namespace detail {
// raw byte transfers of the xvecs / index streams
inline void get_bytes(std::istream &s, void *dst, size_t nbytes) { s.read(static_cast<char *>(dst), (std::streamsize)nbytes); }
inline void put_bytes(std::ostream &s, const void *src, size_t nbytes) { s.write(static_cast<const char *>(src), (std::streamsize)nbytes); }
// the uint32 dimension word in front of every xvecs record; a mismatch ends the program like the reference does
inline void expect_dim(std::istream &s, size_t want)
{
    uint32_t got = 0;
    get_bytes(s, &got, sizeof got);
    if (got != want) {
        std::cout << "file error\n";
        exit(1);
    }
}
}  // namespace detail

template <class P> void read_variable(std::istream &s, P &value) { detail::get_bytes(s, &value, sizeof(P)); }
template <class P> void write_variable(std::ostream &s, const P &value) { detail::put_bytes(s, &value, sizeof(P)); }

// uint32 element count followed by the elements (reference utils.h:45-62)
template <class E> void read_vector(std::istream &s, std::vector<E> &v)
{
    uint32_t count = 0;
    detail::get_bytes(s, &count, sizeof count);
    v.resize(count);
    if (count) detail::get_bytes(s, v.data(), (size_t)count * sizeof(E));
}
template <class E> void write_vector(std::ostream &s, std::vector<E> &v)
{
    const uint32_t count = (uint32_t)v.size();
    detail::put_bytes(s, &count, sizeof count);
    if (count) detail::put_bytes(s, v.data(), (size_t)count * sizeof(E));
}

// nrec records of [uint32 dim][dim x E]; a record whose dim differs from dim aborts (reference utils.h:64-75,91-94)
template <class E> void readXvec(std::ifstream &s, E *rows, const size_t dim, const size_t nrec = 1)
{
    for (size_t r = 0; r < nrec; r++) {
        detail::expect_dim(s, dim);
        detail::get_bytes(s, rows + r * dim, dim * sizeof(E));
    }
}
template <class E> void writeXvec(std::ofstream &s, E *rows, const size_t dim, const size_t nrec = 1)
{
    const uint32_t word = (uint32_t)dim;
    for (size_t r = 0; r < nrec; r++) {
        detail::put_bytes(s, &word, sizeof word);
        detail::put_bytes(s, rows + r * dim, dim * sizeof(E));
    }
}
// like readXvec but widens every component to float (reference utils.h:86-100)
template <class E> void readXvecFvec(std::ifstream &s, float *rows, const size_t dim, const size_t nrec = 1)
{
    std::vector<E> one(dim);
    for (size_t r = 0; r < nrec; r++) {
        detail::expect_dim(s, dim);
        detail::get_bytes(s, one.data(), dim * sizeof(E));
        for (size_t c = 0; c < dim; c++) rows[r * dim + c] = (float)one[c];
    }
}

inline bool exists(const char *path)
{
    struct stat st;
    return path && stat(path, &st) == 0;
}

enum vec_t { base_vec = 0, centroid_vec = 1 };   // which file getL2Distance reads (reference utils.h:134-137)

typedef struct SearchInfo {
    float distance;
    long label;
} SearchInfo_t;
// ascending distance; distances within 0.001 of each other are ordered by label (reference utils.cpp:192-202)
bool cmp(SearchInfo_t a, SearchInfo_t b);
// distance between the query and record vec_id of a .bvecs base file / .fvecs centroid file (utils.cpp:98-137)
float getL2Distance(const float *query, const char *path_vec, const size_t dim, const long vec_id, vec_t type_v);
// records of a .bvecs file of dimension vec_dim (0 when it does not exist; utils.cpp:204-219)
size_t base_vec_num(const char *path_base, size_t vec_dim);
// sorted names of the files of path_dir that end in file_ext; every name must start with file_prefix (utils.cpp:221-272)
void get_files(const char *path_dir, const char *file_ext, std::vector<std::string> &file_list);
void check_files(const char *file_prefix, std::vector<std::string> &file_list);
void get_index_name(const char *path_idx, size_t idx, char *idx_name);   // "<path>_%02lu.index"

// first sub_nx rows of a seed-1234 faiss::rand_perm of the nx input rows (reference utils.cpp:12-19)
void random_subset(const float *x, float *x_out, size_t d, size_t nx, size_t sub_nx);
// hnswlib-order squared L2 distance over floor(d/16)*16 dims (reference utils.cpp:22-51), host helper
float fvec_L2sqr(const float *x, const float *y, size_t d);

}  // namespace ivfhnsw
