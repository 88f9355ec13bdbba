// End-to-end run of the drop-in C++ API on synthetic data, written the way the reference's mains are
// (tests/test_ivfhnsw_sift1b.cpp, tests/test_ivfhnsw_grouping_sift1b.cpp of uniio/ivf-hnsw): files on disk,
// build_quantizer -> train_pq -> assign -> add_batch / add_group -> compute_centroid_norms -> write ->
// read into a fresh index -> one search() call per query -> recall.  Exit code 0 = all checks passed.
//   usage: dropin_smoke <workdir> [adc|grouping] [opq]
#include <cmath>
#include <queue>
#include <random>
#include <unordered_set>

#include <ivf-hnsw/IndexIVF_HNSW.h>
#include <ivf-hnsw/IndexIVF_HNSW_Grouping.h>
#include <ivf-hnsw/Parser.h>
#include <ivf-hnsw/hnswalg.h>

using namespace hnswlib;
using namespace ivfhnsw;

static void fail(const char *what)
{
    std::cerr << "dropin_smoke: FAILED: " << what << std::endl;
    exit(2);
}

int main(int argc, char **argv)
{
    if (argc < 2) { std::cerr << "usage: dropin_smoke <workdir> [adc|grouping] [opq]\n"; return 64; }
    const std::string dir = argv[1];
    const bool grouping = argc > 2 && std::string(argv[2]) == "grouping";
    const bool opq = argc > 3 && std::string(argv[3]) == "opq";
    const size_t d = 32, nc = 256, nb = 40000, nq = 200, nt = 20000, code_size = 8, nsubc = 16, k = 10;

    // ---- synthetic SIFT-like bytes: blobs around nc/4 means
    std::mt19937 rng(1234);
    std::normal_distribution<float> gauss(0.f, 1.f);
    std::vector<float> means(nc / 4 * d);
    for (float &v : means) v = 60.f + 25.f * gauss(rng);
    auto draw = [&](size_t n, std::vector<uint8_t> &out) {
        out.resize(n * d);
        for (size_t i = 0; i < n; i++) {
            const size_t b = rng() % (nc / 4);
            for (size_t j = 0; j < d; j++) {
                const float v = std::round(means[b * d + j] + 9.f * gauss(rng));
                out[i * d + j] = (uint8_t)std::min(255.f, std::max(0.f, v));
            }
        }
    };
    std::vector<uint8_t> base, learn, query;
    draw(nb, base);
    draw(nt, learn);
    draw(nq, query);
    const std::string p_base = dir + "/base.bvecs", p_learn = dir + "/learn.bvecs", p_q = dir + "/query.bvecs",
                      p_cent = dir + "/centroids.fvecs", p_pq = dir + "/pq.pq", p_norm = dir + "/norm.pq",
                      p_opq = dir + "/opq.vt", p_index = dir + "/ivf.index";
    {
        std::ofstream o(p_base, std::ios::binary); writeXvec<uint8_t>(o, base.data(), d, nb);
        std::ofstream l(p_learn, std::ios::binary); writeXvec<uint8_t>(l, learn.data(), d, nt);
        std::ofstream q(p_q, std::ios::binary); writeXvec<uint8_t>(q, query.data(), d, nq);
        // centroids: the first nc learn vectors, jittered (any file works, k-means is not on the path)
        std::vector<float> cent(nc * d);
        for (size_t i = 0; i < nc * d; i++) cent[i] = (float)learn[i] + 0.25f * gauss(rng);
        std::ofstream c(p_cent, std::ios::binary); writeXvec<float>(c, cent.data(), d, nc);
    }

    // ---- exact ground truth on the host
    std::vector<idx_t> gt(nq);
    for (size_t i = 0; i < nq; i++) {
        float best = 1e30f;
        for (size_t j = 0; j < nb; j++) {
            float s = 0;
            for (size_t t = 0; t < d; t++) {
                const float df = (float)query[i * d + t] - (float)base[j * d + t];
                s += df * df;
            }
            if (s < best) { best = s; gt[i] = (idx_t)j; }
        }
    }

    // ---- the reference's flow
    std::vector<float> massQ(nq * d);
    { std::ifstream in(p_q, std::ios::binary); readXvecFvec<uint8_t>(in, massQ.data(), d, nq); }

    IndexIVF_HNSW *index = grouping ? new IndexIVF_HNSW_Grouping(d, nc, code_size, 8, nsubc)
                                    : new IndexIVF_HNSW(d, nc, code_size, 8);
    IndexIVF_HNSW_Grouping *gindex = grouping ? static_cast<IndexIVF_HNSW_Grouping *>(index) : nullptr;
    index->build_quantizer(p_cent.c_str(), "unused.info", "unused.edges", 16, 500);
    index->do_opq = opq;
    {
        std::vector<float> trainvecs(nt * d), subset(nt / 2 * d);
        std::ifstream in(p_learn, std::ios::binary);
        readXvecFvec<uint8_t>(in, trainvecs.data(), d, nt);
        random_subset(trainvecs.data(), subset.data(), d, nt, nt / 2);
        index->train_pq(nt / 2, subset.data());
        faiss::write_ProductQuantizer(index->pq, p_pq.c_str());
        faiss::write_ProductQuantizer(index->norm_pq, p_norm.c_str());
        if (opq) faiss::write_VectorTransform(index->opq_matrix, p_opq.c_str());
    }
    // swap the codebooks behind the index's back, like the reference's mains do when the files exist
    delete index->pq;
    index->pq = faiss::read_ProductQuantizer(p_pq.c_str());
    delete index->norm_pq;
    index->norm_pq = faiss::read_ProductQuantizer(p_norm.c_str());
    if (opq) {
        delete index->opq_matrix;
        index->opq_matrix = dynamic_cast<faiss::LinearTransform *>(faiss::read_VectorTransform(p_opq.c_str()));
        if (!index->opq_matrix) fail("read_VectorTransform did not return a LinearTransform");
    }
    if (index->pq->centroids.size() != d * 256 || index->norm_pq->centroids.size() != 256) fail("PQ file round trip");

    std::vector<float> batch(nb * d);
    { std::ifstream in(p_base, std::ios::binary); readXvecFvec<uint8_t>(in, batch.data(), d, nb); }
    std::vector<idx_t> precomputed_idx(nb), ids(nb);
    index->quantizer->efSearch = 220;
    index->assign(nb, batch.data(), precomputed_idx.data());
    for (size_t i = 0; i < nb; i++) ids[i] = (idx_t)i;
    for (size_t i = 0; i < nb; i++)
        if (precomputed_idx[i] >= nc) fail("assign label out of range");

    if (!grouping) {
        const size_t half = nb / 2;
        index->add_batch(half, batch.data(), ids.data(), precomputed_idx.data());
        index->add_batch(nb - half, batch.data() + half * d, ids.data() + half, precomputed_idx.data() + half);
    } else {
        std::vector<std::vector<float> > data(nc);
        std::vector<std::vector<idx_t> > gids(nc);
        for (size_t i = 0; i < nb; i++) {
            const idx_t g = precomputed_idx[i];
            data[g].insert(data[g].end(), batch.begin() + i * d, batch.begin() + (i + 1) * d);
            gids[g].push_back((idx_t)i);
        }
#pragma omp parallel for
        for (size_t g = 0; g < nc; g++) gindex->add_group(g, gids[g].size(), data[g].data(), gids[g].data());
    }
    index->compute_centroid_norms();
    if (grouping) gindex->compute_inter_centroid_dists();
    index->write(p_index.c_str());
    delete index;

    // ---- fresh index, read everything back from the files
    index = grouping ? new IndexIVF_HNSW_Grouping(d, nc, code_size, 8, nsubc) : new IndexIVF_HNSW(d, nc, code_size, 8);
    gindex = grouping ? static_cast<IndexIVF_HNSW_Grouping *>(index) : nullptr;
    index->build_quantizer(p_cent.c_str(), "unused.info", "unused.edges");
    index->do_opq = opq;
    delete index->pq;
    index->pq = faiss::read_ProductQuantizer(p_pq.c_str());
    delete index->norm_pq;
    index->norm_pq = faiss::read_ProductQuantizer(p_norm.c_str());
    if (opq) index->opq_matrix = dynamic_cast<faiss::LinearTransform *>(faiss::read_VectorTransform(p_opq.c_str()));
    index->read(p_index.c_str());
    if (opq) index->rotate_quantizer();

    index->nprobe = 16;
    index->max_codes = 100000;
    index->quantizer->efSearch = 80;
    if (grouping) gindex->do_pruning = true;

    size_t correct = 0;
    float distances[k];
    long labels[k];
    StopW stopw = StopW();
    for (size_t i = 0; i < nq; i++) {
        index->search(k, massQ.data() + i * d, distances, labels);
        // faiss max-heap order: the root is the worst of the k
        for (size_t j = 1; j < k; j++)
            if (distances[j] > distances[0]) fail("result is not a max-heap");
        for (size_t j = 0; j < k; j++)
            if (labels[j] == (long)gt[i]) { correct++; break; }
    }
    const float us = stopw.getElapsedTimeMicro() / nq;
    // the batched extension must agree with the one-by-one loop
    std::vector<float> D(nq * k);
    std::vector<long> L(nq * k);
    index->search_batch(nq, k, massQ.data(), D.data(), L.data(), false);
    index->search(k, massQ.data() + (nq - 1) * d, distances, labels);
    for (size_t j = 0; j < k; j++)
        if (L[(nq - 1) * k + j] != labels[j] || D[(nq - 1) * k + j] != distances[j]) fail("search_batch != search");

    const float recall = 1.0f * correct / nq;
    std::cout << "Recall@" << k << ": " << recall << std::endl;
    std::cout << "Time per query: " << us << " us" << std::endl;
    if (recall < 0.5f) fail("recall too low");
    delete index;
    std::cout << "dropin_smoke OK" << std::endl;
    return 0;
}
