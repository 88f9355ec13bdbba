// Host side of the drop-in: ivfhnsw::IndexIVF_HNSW / IndexIVF_HNSW_Grouping (include/ivf-hnsw/*.h) and the
// faiss file helpers (include/faiss/index_io.h), all forwarding to the C ABI of libivfhnsw_b200.so.
// Compiled into libivfhnsw_dropin.so.  No arithmetic of the search / indexing path happens here.
#include <dirent.h>

#include <algorithm>
#include <cstdarg>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <stdexcept>
#include <string>

#include <ivf-hnsw/IndexIVF_HNSW.h>
#include <ivf-hnsw/IndexIVF_HNSW_Grouping.h>

#include "../../include/ivfhnsw_b200.h"

// ------------------------------------------------------------------------------------------ faiss helpers
namespace faiss {

void LinearTransform::apply_noalloc(long n, const float *x, float *xt) const
{
    for (long i = 0; i < n; i++)
        for (int o = 0; o < d_out; o++) {
            float acc = have_bias ? b[o] : 0.f;
            const float *row = A.data() + (size_t)o * d_in;
            for (int j = 0; j < d_in; j++) acc += row[j] * x[(size_t)i * d_in + j];
            xt[(size_t)i * d_out + o] = acc;
        }
}

void LinearTransform::transform_transpose(long n, const float *y, float *x) const
{
    for (long i = 0; i < n; i++)
        for (int j = 0; j < d_in; j++) {
            float acc = 0.f;
            for (int o = 0; o < d_out; o++) {
                const float v = y[(size_t)i * d_out + o] - (have_bias ? b[o] : 0.f);
                acc += A[(size_t)o * d_in + j] * v;
            }
            x[(size_t)i * d_in + j] = acc;
        }
}

namespace {
struct File {
    FILE *f;
    File(const char *name, const char *mode) : f(fopen(name, mode))
    {
        if (!f) throw std::runtime_error(std::string("could not open ") + name);
    }
    ~File() { if (f) fclose(f); }
    void put(const void *p, size_t n) { if (fwrite(p, 1, n, f) != n) throw std::runtime_error("write error"); }
    void get(void *p, size_t n) { if (fread(p, 1, n, f) != n) throw std::runtime_error("read error"); }
    template <typename T> void put1(const T &v) { put(&v, sizeof(T)); }
    template <typename T> T get1() { T v; get(&v, sizeof(T)); return v; }
    void put_floats(const std::vector<float> &v) { put1<size_t>(v.size()); put(v.data(), v.size() * 4); }
    void get_floats(std::vector<float> &v) { v.resize(get1<size_t>()); get(v.data(), v.size() * 4); }
};
}  // namespace

// size_t d, M, nbits; size_t count + floats (faiss/index_io.cpp:117-124,504-513)
void write_ProductQuantizer(const ProductQuantizer *pq, const char *fname)
{
    File f(fname, "wb");
    f.put1<size_t>(pq->d);
    f.put1<size_t>(pq->M);
    f.put1<size_t>(pq->nbits);
    f.put_floats(pq->centroids);
}

ProductQuantizer *read_ProductQuantizer(const char *fname)
{
    File f(fname, "rb");
    const size_t d = f.get1<size_t>(), M = f.get1<size_t>(), nbits = f.get1<size_t>();
    ProductQuantizer *pq = new ProductQuantizer(d, M, nbits);
    f.get_floats(pq->centroids);
    return pq;
}

// "LTra", bool have_bias, A, b, int d_in, int d_out, bool is_trained (faiss/index_io.cpp:63-105,425-470)
void write_VectorTransform(const VectorTransform *vt, const char *fname)
{
    const LinearTransform *lt = dynamic_cast<const LinearTransform *>(vt);
    if (!lt) throw std::runtime_error("only LinearTransform can be written");
    File f(fname, "wb");
    f.put("LTra", 4);
    f.put1<bool>(lt->have_bias);
    f.put_floats(lt->A);
    f.put_floats(lt->b);
    f.put1<int>(lt->d_in);
    f.put1<int>(lt->d_out);
    f.put1<bool>(lt->is_trained);
}

VectorTransform *read_VectorTransform(const char *fname)
{
    File f(fname, "rb");
    char tag[4];
    f.get(tag, 4);
    // OPQ matrices are stored either as generic linear transforms or with their own tag plus (M, niter, ...)
    const bool opq = !memcmp(tag, "OPQm", 4), rrot = !memcmp(tag, "rrot", 4);
    if (memcmp(tag, "LTra", 4) && !opq && !rrot) throw std::runtime_error("unsupported VectorTransform file");
    LinearTransform *lt = new LinearTransform();
    lt->have_bias = f.get1<bool>();
    f.get_floats(lt->A);
    f.get_floats(lt->b);
    lt->d_in = f.get1<int>();
    lt->d_out = f.get1<int>();
    lt->is_trained = f.get1<bool>();
    return lt;
}

}  // namespace faiss

// ------------------------------------------------------------------------------------------ utils
namespace ivfhnsw {

namespace {
// faiss::RandomGenerator on linux + rand_perm (faiss/utils.cpp:139-159,311-321): glibc random_r, 8-byte state
void faiss_rand_perm(std::vector<int> &perm, size_t n, long seed)
{
    char state[8];
    struct random_data rd;
    memset(&rd, 0, sizeof rd);
    initstate_r((unsigned)seed, state, sizeof state, &rd);
    perm.resize(n);
    for (size_t i = 0; i < n; i++) perm[i] = (int)i;
    for (size_t i = 0; i + 1 < n; i++) {
        int32_t r;
        random_r(&rd, &r);
        std::swap(perm[i], perm[i + (size_t)(r % (int)(n - i))]);
    }
}

uint64_t fingerprint(const float *p, size_t n)
{
    // pointer + size + a strided sample of the contents: cheap enough for the one-query search() path
    uint64_t h = 1469598103934665603ull ^ (uint64_t)(uintptr_t)p;
    h = (h ^ n) * 1099511628211ull;
    const size_t step = n > 256 ? n / 256 : 1;
    for (size_t i = 0; i < n; i += step) {
        uint32_t b;
        memcpy(&b, p + i, 4);
        h = (h ^ b) * 1099511628211ull;
    }
    if (n) {
        uint32_t b;
        memcpy(&b, p + n - 1, 4);
        h = (h ^ b) * 1099511628211ull;
    }
    return h;
}

void to_max_heap_order(size_t k, float *d, long *l)
{
    // ascending -> descending: a descending array is a valid faiss max-heap (root = worst), which is what
    // reference callers that peek at distances[0] expect (faiss/Heap.h:89-143)
    std::reverse(d, d + k);
    std::reverse(l, l + k);
}
}  // namespace

void random_subset(const float *x, float *x_out, size_t d, size_t nx, size_t sub_nx)
{
    std::vector<int> perm;
    faiss_rand_perm(perm, nx, 1234);
    for (size_t i = 0; i < sub_nx; i++) memcpy(x_out + i * d, x + (size_t)perm[i] * d, sizeof(float) * d);
}

bool cmp(SearchInfo_t a, SearchInfo_t b)
{
    if (a.distance > b.distance) return false;
    if (std::abs(b.distance - a.distance) <= 0.001f) return a.label < b.label;   // "equal" distances: by label
    return true;
}

float getL2Distance(const float *query, const char *path_vec, const size_t dim, const long vec_id, vec_t type_v)
{
    std::vector<float> v(dim);
    std::ifstream in(path_vec, std::ios::binary);
    if (type_v == base_vec) {          // .bvecs record: [uint32 dim][dim x uint8]
        in.seekg((std::streamoff)vec_id * (std::streamoff)(sizeof(uint32_t) + dim), in.beg);
        readXvecFvec<uint8_t>(in, v.data(), dim, 1);
    } else if (type_v == centroid_vec) {   // .fvecs record: [uint32 dim][dim x float]
        in.seekg((std::streamoff)vec_id * (std::streamoff)(sizeof(uint32_t) + dim * sizeof(float)), in.beg);
        readXvec<float>(in, v.data(), dim, 1);
    } else {
        std::cout << "Invalid vector type: " << type_v << std::endl;
        assert(0);
    }
    return fvec_L2sqr(query, v.data(), dim);
}

size_t base_vec_num(const char *path_base, size_t vec_dim)
{
    struct stat st;
    if (stat(path_base, &st)) return 0;
    const size_t rec = sizeof(uint32_t) + vec_dim * sizeof(uint8_t);
    const size_t n = (size_t)st.st_size / rec;
    if ((size_t)st.st_size != rec * n) {
        std::cout << "Invalid size of file: " << path_base << std::endl;
        assert(0);
    }
    return n;
}

void get_files(const char *path_dir, const char *file_ext, std::vector<std::string> &file_list)
{
    DIR *dir = opendir(path_dir);
    if (!dir) {
        std::cout << "Failed to open dir: " << path_dir << std::endl;
        return;
    }
    const size_t le = strlen(file_ext);
    while (struct dirent *e = readdir(dir)) {
        const size_t ln = strlen(e->d_name);
        // a name that ends in the extension and is longer than it
        if (ln > le && !strcmp(e->d_name + ln - le, file_ext)) file_list.push_back(e->d_name);
    }
    closedir(dir);
    std::sort(file_list.begin(), file_list.end());
}

void check_files(const char *file_prefix, std::vector<std::string> &file_list)
{
    for (const std::string &f : file_list)
        if (f.compare(0, strlen(file_prefix), file_prefix) != 0) assert(0);
}

void get_index_name(const char *path_idx, size_t idx, char *idx_name)
{
    sprintf(idx_name, "%s_%02lu%s", path_idx, (unsigned long)idx, ".index");
}

float fvec_L2sqr(const float *x, const float *y, size_t d)
{
    float lane[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const size_t n = (d >> 4) << 4;
    for (size_t i = 0; i < n; i += 8)
        for (size_t j = 0; j < 8; j++) {
            const float df = x[i + j] - y[i + j];
            lane[j] += df * df;
        }
    float r = lane[0];
    for (int j = 1; j < 8; j++) r += lane[j];
    return r;
}

// ------------------------------------------------------------------------------------------ IndexIVF_HNSW
void IndexIVF_HNSW::ck(int rc, const char *what) const
{
    if (rc >= 0) return;
    fprintf(stderr, "ivfhnsw_b200: %s failed: %s\n", what, h_ ? ivfhnsw_b200_last_error(h_) : "no handle");
    exit(1);
}

void IndexIVF_HNSW::init(size_t dim, size_t ncentroids, size_t bytes_per_code, size_t nbits_per_idx, size_t nsubc)
{
    d = dim;
    nc = ncentroids;
    code_size = bytes_per_code;
    nsubc_ = nsubc;
    quantizer = nullptr;
    pq = new faiss::ProductQuantizer(dim, bytes_per_code, nbits_per_idx);
    norm_pq = new faiss::ProductQuantizer(1, 1, nbits_per_idx);
    opq_matrix = nullptr;
    do_opq = false;
    nprobe = 16;
    max_codes = 10000;
    pruning_flag_ = false;
    M = 16;
    memset(&hdr_idx, 0, sizeof hdr_idx);
    fp_pq_ = fp_norm_ = fp_opq_ = 0;
    sent_nprobe_ = sent_max_codes_ = (size_t)-1;
    sent_flags_ = -1;
    h_ = nullptr;
    const char *dev = getenv("IVFHNSW_B200_DEVICE");
    const int rc = ivfhnsw_b200_create(&h_, dim, ncentroids, bytes_per_code, nbits_per_idx, nsubc, dev ? atoi(dev) : 0);
    if (rc < 0) {
        fprintf(stderr, "ivfhnsw_b200: create failed: %s\n", h_ ? ivfhnsw_b200_last_error(h_) : "no sm_100 device");
        exit(1);
    }
}

IndexIVF_HNSW::IndexIVF_HNSW(size_t dim, size_t ncentroids, size_t bytes_per_code, size_t nbits_per_idx, size_t)
{
    init(dim, ncentroids, bytes_per_code, nbits_per_idx, 0);
}

IndexIVF_HNSW::IndexIVF_HNSW(size_t dim, size_t ncentroids, size_t bytes_per_code, size_t nbits_per_idx, size_t nsubc,
                             int)
{
    init(dim, ncentroids, bytes_per_code, nbits_per_idx, nsubc);
}

IndexIVF_HNSW::~IndexIVF_HNSW()
{
    if (quantizer) delete quantizer;
    if (pq) delete pq;
    if (norm_pq) delete norm_pq;
    if (opq_matrix) delete opq_matrix;
    if (h_) ivfhnsw_b200_destroy(h_);
}

std::vector<float> &IndexIVF_HNSW::get_centroid_norms()
{
    centroid_norms.resize(nc);
    ck(ivfhnsw_b200_get_centroid_norms(h_, centroid_norms.data()), "get_centroid_norms");
    return centroid_norms;
}

void IndexIVF_HNSW::build_quantizer(const char *path_data, const char *, const char *, size_t M, size_t efConstruction)
{
    this->M = M;
    hdr_idx.efConstruction = (uint32_t)efConstruction;   // reference IndexIVF_HNSW.cpp:43
    // only the centroid table is needed: the coarse quantizer is the exact GEMM over it, no graph files
    std::vector<float> cent(nc * d);
    {
        std::ifstream in(path_data, std::ios::binary);
        if (!in) {
            fprintf(stderr, "ivfhnsw_b200: cannot open centroid file %s\n", path_data);
            exit(1);
        }
        readXvec<float>(in, cent.data(), d, nc);
    }
    ck(ivfhnsw_b200_set_centroids(h_, cent.data()), "set_centroids");
    if (quantizer) delete quantizer;
    quantizer = new hnswlib::HierarchicalNSW(d, nc, M, efConstruction);
}

void IndexIVF_HNSW::sync_quantizers()
{
    if (pq) {
        if (pq->centroids.size() != d * 256) {
            fprintf(stderr, "ivfhnsw_b200: residual PQ has %zu floats, expected %zu\n", pq->centroids.size(), d * 256);
            exit(1);
        }
        const uint64_t fp = fingerprint(pq->centroids.data(), pq->centroids.size());
        if (fp != fp_pq_) {
            ck(ivfhnsw_b200_set_pq(h_, pq->centroids.data()), "set_pq");
            fp_pq_ = fp;
        }
    }
    if (norm_pq && norm_pq->centroids.size() == 256) {
        const uint64_t fp = fingerprint(norm_pq->centroids.data(), 256);
        if (fp != fp_norm_) {
            ck(ivfhnsw_b200_set_norm_pq(h_, norm_pq->centroids.data()), "set_norm_pq");
            fp_norm_ = fp;
        }
    }
    if (opq_matrix && opq_matrix->A.size() == d * d) {
        const uint64_t fp = fingerprint(opq_matrix->A.data(), d * d);
        if (fp != fp_opq_) {
            ck(ivfhnsw_b200_set_opq(h_, opq_matrix->A.data()), "set_opq");
            fp_opq_ = fp;
        }
    }
    const int flags = (do_opq ? 1 : 0) | (pruning_flag_ ? 2 : 0);
    if (nprobe != sent_nprobe_ || max_codes != sent_max_codes_ || flags != sent_flags_) {
        ck(ivfhnsw_b200_set_params(h_, nprobe, max_codes, do_opq ? 1 : 0, pruning_flag_ ? 1 : 0), "set_params");
        sent_nprobe_ = nprobe;
        sent_max_codes_ = max_codes;
        sent_flags_ = flags;
    }
}

void IndexIVF_HNSW::pull_quantizers()
{
    pq->centroids.resize(d * 256);
    norm_pq->centroids.resize(256);
    std::vector<float> A(do_opq ? d * d : 0);
    ck(ivfhnsw_b200_get_pq(h_, pq->centroids.data(), norm_pq->centroids.data(), do_opq ? A.data() : nullptr), "get_pq");
    fp_pq_ = fingerprint(pq->centroids.data(), pq->centroids.size());
    fp_norm_ = fingerprint(norm_pq->centroids.data(), 256);
    if (do_opq) {
        if (opq_matrix) delete opq_matrix;
        opq_matrix = new faiss::LinearTransform((int)d, (int)d, false);
        opq_matrix->A.swap(A);
        fp_opq_ = fingerprint(opq_matrix->A.data(), d * d);
    }
}

void IndexIVF_HNSW::assign(size_t n, const float *x, idx_t *labels, size_t k)
{
    // the reference returns the k-th nearest for k > 1 (top of a k-element max-heap, IndexIVF_HNSW.cpp:71) and so does
    // the C ABI; every caller passes k = 1
    ck(ivfhnsw_b200_assign(h_, n, x, labels, k), "assign");
}

void IndexIVF_HNSW::assign_u8(size_t n, const uint8_t *x, idx_t *labels)
{
    ck(ivfhnsw_b200_assign_u8(h_, n, x, labels), "assign_u8");
}

void IndexIVF_HNSW::search_batch(size_t nq, size_t k, const float *x, float *distances, long *labels, bool sorted)
{
    static_assert(sizeof(long) == sizeof(int64_t), "labels are 64-bit");
    sync_quantizers();
    ck(ivfhnsw_b200_search(h_, nq, k, x, distances, reinterpret_cast<int64_t *>(labels)), "search");
    if (!sorted)
        for (size_t i = 0; i < nq; i++) to_max_heap_order(k, distances + i * k, labels + i * k);
}

void IndexIVF_HNSW::search(size_t k, const float *x, float *distances, long *labels)
{
    last_query_.assign(x, x + d);
    search_batch(1, k, x, distances, labels, false);
}

// Tracing of the probed centroids (reference IndexIVF_HNSW.cpp:298-326).  The reference records them inside every
// search(); here they are recomputed on demand for the last query (one coarse call) so that the hot path stays clean.
void IndexIVF_HNSW::trace_centroids(size_t idx_q, bool missed)
{
    trace_query_centroid_dists.clear();
    trace_centroid_idxs.clear();
    if (last_query_.size() == d) {
        std::vector<float> cd(nprobe);
        std::vector<idx_t> ci(nprobe);
        sync_quantizers();
        ck(ivfhnsw_b200_coarse(h_, 1, last_query_.data(), nprobe, cd.data(), ci.data()), "coarse");
        for (size_t j = nprobe; j-- > 0;) {     // farthest first, like the unload loop of IndexIVF_HNSW.cpp:249-256
            trace_query_centroid_dists.push_back(cd[j]);
            trace_centroid_idxs.push_back(ci[j]);
        }
    }
    std::ofstream log_trace(missed ? "centroids_miss.log" : "centroids_hit.log", std::ofstream::out | std::ofstream::app);
    if (!log_trace.is_open()) {
        std::cout << "Failed to open log file for centroids traceing" << std::endl;
        return;
    }
    std::vector<uint32_t> sizes(nc);
    ck(ivfhnsw_b200_list_sizes(h_, sizes.data()), "list_sizes");
    const size_t sz = trace_centroid_idxs.size();
    std::cout << "centroids number " << sz << std::endl;
    if (sz > 0) log_trace << "query " << idx_q << " centroids info\n";
    for (size_t i = 0; i < sz; i++)
        log_trace << "centroid " << trace_centroid_idxs[i] << " with distance " << trace_query_centroid_dists[i]
                  << " with group size " << sizes[trace_centroid_idxs[i]] << "\n";
}

void IndexIVF_HNSW::add_batch2(size_t n, const float *x, const idx_t *xids, const idx_t *idx, uint64_t *eids, char *obuf)
{
    sync_quantizers();
    std::vector<uint8_t> xcodes(n * code_size), xnorm(n);
    ck(ivfhnsw_b200_add_batch_codes(h_, n, x, xids, idx, xcodes.data(), xnorm.data()), "add_batch2");
    char *p = obuf;     // record = [uint64 external id][code_size code bytes][norm code], IndexIVF_HNSW.cpp:181-203
    for (size_t i = 0; i < n; i++) {
        memcpy(p, eids + i, sizeof(uint64_t));
        p += sizeof(uint64_t);
        memcpy(p, xcodes.data() + i * code_size, code_size);
        p += code_size;
        *p++ = (char)xnorm[i];
    }
}

int IndexIVF_HNSW::copy_file(const char *file_src, const char *file_dst)
{
    std::ifstream in(file_src, std::ios::binary);
    if (!in) return -1;
    std::ofstream out(file_dst, std::ios::binary);
    out << in.rdbuf();
    return out.good() ? 0 : -1;
}

// ORCV directory (reference IndexIVF_HNSW.cpp:672-755).  Not reproduced: the reference's write2 also rotates the
// quantizer's centroids IN PLACE while producing opq.vec ("TODO: confirm it" there); the file holds what it holds in
// the reference -- one d-vector, the last centroid -- and the centroid table is left alone.
void IndexIVF_HNSW::write2(const char *home_dir, size_t n_vecs, bool do_opq_flag, const char *path_edge)
{
    sync_host_lists();
    char path_f[512];
    {
        snprintf(path_f, sizeof path_f, "%s/hdr.vec", home_dir);
        std::cout << "write file: " << path_f << std::endl;
        std::ofstream output(path_f, std::ios::binary);
        hdr_idx.n = (uint32_t)n_vecs;
        hdr_idx.nc = (uint32_t)nc;
        hdr_idx.code_size = (uint32_t)code_size;
        hdr_idx.code_bytes = (uint32_t)(code_size / 8);
        hdr_idx.d = (uint32_t)d;
        hdr_idx.do_opq = do_opq_flag ? 1 : 0;
        hdr_idx.M = (uint32_t)M;
        hdr_idx.dmatch = dmatch;
        hdr_idx.dnear = dnear;
        write_variable(output, hdr_idx);
    }
    {
        snprintf(path_f, sizeof path_f, "%s/cnorms.vec", home_dir);
        std::cout << "write file: " << path_f << std::endl;
        std::ofstream output(path_f, std::ios::binary);
        write_vector(output, get_centroid_norms());
    }
    {
        snprintf(path_f, sizeof path_f, "%s/pq.vec", home_dir);
        std::ofstream output(path_f, std::ios::binary);
        for (size_t i = 0; i < nc; i++) write_vector(output, codes[i]);
    }
    if (do_opq_flag) {
        std::vector<float> cent(nc * d), last(d);
        ck(ivfhnsw_b200_get_centroids(h_, cent.data()), "get_centroids");
        memcpy(last.data(), cent.data() + (nc - 1) * d, d * sizeof(float));
        snprintf(path_f, sizeof path_f, "%s/opq.vec", home_dir);
        std::ofstream output(path_f, std::ios::binary);
        write_vector(output, last);
    }
    {
        snprintf(path_f, sizeof path_f, "%s/normpq.vec", home_dir);
        std::ofstream output(path_f, std::ios::binary);
        for (size_t i = 0; i < nc; i++) write_vector(output, norm_codes[i]);
    }
    snprintf(path_f, sizeof path_f, "%s/edge.vec", home_dir);
    copy_file(path_edge, path_f);
}

IndexIVF_HNSW::idx_t IndexIVF_HNSW::search_enn(const float *x, float *distances, long *labels)
{
    search_batch(1, 1, x, distances, labels, true);
    return labels[0] < 0 ? 0 : 1;
}

void IndexIVF_HNSW::search2(size_t k, const float *x, float *distances, long *labels, float *query_centroid_dists,
                            idx_t *centroid_idxs)
{
    sync_quantizers();
    ck(ivfhnsw_b200_search_given_coarse(h_, 1, k, x, query_centroid_dists, centroid_idxs, distances,
                                        reinterpret_cast<int64_t *>(labels)),
       "search2");
    to_max_heap_order(k, distances, labels);
}

void IndexIVF_HNSW::add_batch(size_t n, const float *x, const idx_t *xids, const idx_t *precomputed_idx)
{
    sync_quantizers();
    ck(ivfhnsw_b200_add_batch(h_, n, x, xids, precomputed_idx), "add_batch");
}

void IndexIVF_HNSW::train_pq(size_t n, const float *x)
{
    // train_pq replaces all codebooks: push only the parameters (do_opq), not the stale codebooks
    ck(ivfhnsw_b200_set_params(h_, nprobe, max_codes, do_opq ? 1 : 0, pruning_flag_ ? 1 : 0), "set_params");
    sent_flags_ = -1;
    ck(ivfhnsw_b200_train_pq(h_, n, x), "train_pq");
    pull_quantizers();
}

void IndexIVF_HNSW::compute_centroid_norms()
{
    ck(ivfhnsw_b200_compute_centroid_norms(h_), "compute_centroid_norms");
}

void IndexIVF_HNSW::rotate_quantizer()
{
    if (!do_opq) {   // reference IndexIVF_HNSW.cpp:790-793
        printf("OPQ encoding is turned off\n");
        abort();
    }
    sync_quantizers();
    ck(ivfhnsw_b200_rotate_quantizer(h_), "rotate_quantizer");
}

namespace {
struct HostLists {
    std::vector<int64_t> off;
    std::vector<uint32_t> ids, nn, ss;
    std::vector<uint8_t> codes, ncodes;
    std::vector<float> alphas, icd;
};
}  // namespace

void IndexIVF_HNSW::sync_host_lists()
{
    ck(ivfhnsw_b200_finalize(h_), "finalize");
    size_t n = 0;
    ck(ivfhnsw_b200_ntotal(h_, &n), "ntotal");
    HostLists L;
    L.off.resize(nc + 1);
    L.ids.resize(std::max<size_t>(n, 1));
    L.codes.resize(std::max<size_t>(n, 1) * code_size);
    L.ncodes.resize(std::max<size_t>(n, 1));
    if (nsubc_) { L.nn.resize(nc * nsubc_); L.ss.resize(nc * nsubc_); L.alphas.resize(nc); L.icd.resize(nc * nsubc_); }
    ck(ivfhnsw_b200_export_lists(h_, L.off.data(), L.ids.data(), L.codes.data(), L.ncodes.data(),
                                 nsubc_ ? L.nn.data() : nullptr, nsubc_ ? L.ss.data() : nullptr,
                                 nsubc_ ? L.alphas.data() : nullptr, nsubc_ ? L.icd.data() : nullptr),
       "export_lists");
    ids.assign(nc, {});
    codes.assign(nc, {});
    norm_codes.assign(nc, {});
    for (size_t c = 0; c < nc; c++) {
        const size_t a = (size_t)L.off[c], b = (size_t)L.off[c + 1];
        ids[c].assign(L.ids.begin() + a, L.ids.begin() + b);
        codes[c].assign(L.codes.begin() + a * code_size, L.codes.begin() + b * code_size);
        norm_codes[c].assign(L.ncodes.begin() + a, L.ncodes.begin() + b);
    }
}

// File layout (reference IndexIVF_HNSW.cpp:637-663,758-779; Grouping.cpp:397-483): written / read natively by the
// C ABI (ivfhnsw_b200_write_index / ivfhnsw_b200_read_index, csrc/io.cu) straight from / into the device lists.
void IndexIVF_HNSW::write(const char *path_index)
{
    ck(ivfhnsw_b200_write_index(h_, path_index), "write");
}
void IndexIVF_HNSW::write(const char *path_index, bool) { write(path_index); }

void IndexIVF_HNSW::read(const char *path)
{
    sync_quantizers();
    ck(ivfhnsw_b200_read_index(h_, path), "read");
}

// ------------------------------------------------------------------------------------------ Grouping
IndexIVF_HNSW_Grouping::IndexIVF_HNSW_Grouping(size_t dim, size_t ncentroids, size_t bytes_per_code,
                                               size_t nbits_per_idx, size_t nsubcentroids)
    : IndexIVF_HNSW(dim, ncentroids, bytes_per_code, nbits_per_idx, nsubcentroids, 0), nsubc(nsubcentroids),
      do_pruning(false)
{
}

void IndexIVF_HNSW_Grouping::add_group(size_t group_idx, size_t group_size, const float *x, const idx_t *gids)
{
    // the C ABI serialises callers on the handle's mutex; quantizer sync must not race either
    pruning_flag_ = do_pruning;
    {
        static std::mutex mu;   // only guards the (rare) re-upload of swapped codebooks
        std::lock_guard<std::mutex> lk(mu);
        sync_quantizers();
    }
    ck(ivfhnsw_b200_add_group(h_, group_idx, group_size, x, gids), "add_group");
}

void IndexIVF_HNSW_Grouping::search(size_t k, const float *x, float *distances, long *labels)
{
    pruning_flag_ = do_pruning;
    IndexIVF_HNSW::search(k, x, distances, labels);
}

void IndexIVF_HNSW_Grouping::searchDisk_batch(size_t nq, size_t k, const float *queries, float *distances, long *labels,
                                              const char *path_base)
{
    // 2k candidates from the index (the reference's intent, IndexIVF_HNSW_Grouping.cpp:373-374), ascending
    const size_t k2 = 2 * k;
    std::vector<float> d2(nq * k2);
    std::vector<long> l2(nq * k2);
    pruning_flag_ = do_pruning;
    search_batch(nq, k2, queries, d2.data(), l2.data(), true);
    // their raw vectors from the .bvecs file (getL2Distance -> readBaseVec, utils.cpp:98-107), expanded to float
    std::vector<float> cand(nq * k2 * d, 0.f);
    std::ifstream in(path_base, std::ios::binary);
    if (!in) {
        fprintf(stderr, "ivfhnsw_b200: cannot open base file %s\n", path_base);
        exit(1);
    }
    for (size_t i = 0; i < nq * k2; i++) {
        if (l2[i] < 0) continue;
        in.seekg((std::streamoff)l2[i] * (std::streamoff)(sizeof(uint32_t) + d), in.beg);
        readXvecFvec<uint8_t>(in, cand.data() + i * d, d, 1);
    }
    static_assert(sizeof(long) == sizeof(int64_t), "labels are 64-bit");
    ck(ivfhnsw_b200_rerank(h_, nq, k2, k, queries, cand.data(), reinterpret_cast<const int64_t *>(l2.data()), distances,
                           reinterpret_cast<int64_t *>(labels)),
       "rerank");
}

void IndexIVF_HNSW_Grouping::searchDisk(size_t k, const float *query, float *distances, long *labels, const char *path_base)
{
    searchDisk_batch(1, k, query, distances, labels, path_base);
}

void IndexIVF_HNSW_Grouping::dump_inter_centroid_dists(char *path)
{
    FILE *fp = fopen(path, "w");
    if (!fp) {
        std::cout << "Failed to open file: " << path << std::endl;
        return;
    }
    ck(ivfhnsw_b200_finalize(h_), "finalize");
    size_t n = 0;
    ck(ivfhnsw_b200_ntotal(h_, &n), "ntotal");
    std::vector<int64_t> off(nc + 1);
    std::vector<uint32_t> ids_(std::max<size_t>(n, 1)), nn(nc * nsubc), ss(nc * nsubc);
    std::vector<uint8_t> cd(std::max<size_t>(n, 1) * code_size), ncd(std::max<size_t>(n, 1));
    std::vector<float> icd(nc * nsubc), al(nc);
    ck(ivfhnsw_b200_export_lists(h_, off.data(), ids_.data(), cd.data(), ncd.data(), nn.data(), ss.data(), al.data(),
                                 icd.data()),
       "export_lists");
    for (size_t i = 0; i < nc; i++)
        for (size_t subc = 0; subc < nsubc; subc++)
            fprintf(fp, "distance of centriod %lu to centriod %lu is %f\n", (unsigned long)i, (unsigned long)subc,
                    icd[i * nsubc + subc]);
    fclose(fp);
}

void IndexIVF_HNSW_Grouping::train_pq(size_t n, const float *x)
{
    pruning_flag_ = do_pruning;
    IndexIVF_HNSW::train_pq(n, x);
}

void IndexIVF_HNSW_Grouping::compute_inter_centroid_dists()
{
    ck(ivfhnsw_b200_finalize(h_), "finalize");   // neighbour lists, inter-centroid distances, alphas, CSR
}

void IndexIVF_HNSW_Grouping::write(const char *path_index)
{
    ck(ivfhnsw_b200_write_index(h_, path_index), "write");
}
void IndexIVF_HNSW_Grouping::write(const char *path_index, bool) { write(path_index); }

void IndexIVF_HNSW_Grouping::read(const char *path_index)
{
    pruning_flag_ = do_pruning;
    sync_quantizers();
    ck(ivfhnsw_b200_read_index(h_, path_index), "read");
}

void IndexIVF_HNSW_Grouping::sync_host_lists()
{
    IndexIVF_HNSW::sync_host_lists();
    std::vector<int64_t> off(nc + 1);
    size_t n = 0;
    ck(ivfhnsw_b200_ntotal(h_, &n), "ntotal");
    std::vector<uint32_t> ids_(std::max<size_t>(n, 1)), nn(nc * nsubc), ss(nc * nsubc);
    std::vector<uint8_t> cd(std::max<size_t>(n, 1) * code_size), ncd(std::max<size_t>(n, 1));
    std::vector<float> icd(nc * nsubc);
    alphas.resize(nc);
    ck(ivfhnsw_b200_export_lists(h_, off.data(), ids_.data(), cd.data(), ncd.data(), nn.data(), ss.data(), alphas.data(),
                                 icd.data()),
       "export_lists");
    nn_centroid_idxs.assign(nc, {});
    subgroup_sizes.assign(nc, {});
    for (size_t c = 0; c < nc; c++) {
        nn_centroid_idxs[c].assign(nn.begin() + c * nsubc, nn.begin() + (c + 1) * nsubc);
        subgroup_sizes[c].assign(ss.begin() + c * nsubc, ss.begin() + (c + 1) * nsubc);
    }
}

}  // namespace ivfhnsw
