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

template <typename T> void read_variable(std::istream &in, T &pod) { in.read(reinterpret_cast<char *>(&pod), sizeof(T)); }
template <typename T> void write_variable(std::ostream &out, const T &pod) { out.write(reinterpret_cast<const char *>(&pod), sizeof(T)); }

// uint32 element count followed by the elements (reference utils.h:45-62)
template <typename T> void read_vector(std::istream &in, std::vector<T> &vec)
{
    uint32_t size = 0;
    in.read(reinterpret_cast<char *>(&size), sizeof(uint32_t));
    vec.resize(size);
    if (size) in.read(reinterpret_cast<char *>(vec.data()), (std::streamsize)size * sizeof(T));
}
template <typename T> void write_vector(std::ostream &out, std::vector<T> &vec)
{
    const uint32_t size = (uint32_t)vec.size();
    out.write(reinterpret_cast<const char *>(&size), sizeof(uint32_t));
    if (size) out.write(reinterpret_cast<const char *>(vec.data()), (std::streamsize)size * sizeof(T));
}

// n records of [uint32 dim][dim x T]; a record whose dim differs from d aborts (reference utils.h:64-75,91-94)
template <typename T> void readXvec(std::ifstream &in, T *data, const size_t d, const size_t n = 1)
{
    for (size_t i = 0; i < n; i++) {
        uint32_t dim = 0;
        in.read(reinterpret_cast<char *>(&dim), sizeof(uint32_t));
        if (dim != d) {
            std::cout << "file error\n";
            exit(1);
        }
        in.read(reinterpret_cast<char *>(data + i * d), (std::streamsize)(d * sizeof(T)));
    }
}
template <typename T> void writeXvec(std::ofstream &out, T *data, const size_t d, const size_t n = 1)
{
    const uint32_t dim = (uint32_t)d;
    for (size_t i = 0; i < n; i++) {
        out.write(reinterpret_cast<const char *>(&dim), sizeof(uint32_t));
        out.write(reinterpret_cast<const char *>(data + i * d), (std::streamsize)(d * sizeof(T)));
    }
}
// like readXvec but widens every component to float (reference utils.h:86-100)
template <typename T> void readXvecFvec(std::ifstream &in, float *data, const size_t d, const size_t n = 1)
{
    std::vector<T> row(d);
    for (size_t i = 0; i < n; i++) {
        uint32_t dim = 0;
        in.read(reinterpret_cast<char *>(&dim), sizeof(uint32_t));
        if (dim != d) {
            std::cout << "file error\n";
            exit(1);
        }
        in.read(reinterpret_cast<char *>(row.data()), (std::streamsize)(d * sizeof(T)));
        for (size_t j = 0; j < d; j++) data[i * d + j] = (float)row[j];
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
