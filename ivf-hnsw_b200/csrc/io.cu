// Index files in the reference's own format, host side of the C ABI:
//   IndexIVF_HNSW::write / read                     (IndexIVF_HNSW.cpp:637-663, 758-779)
//   IndexIVF_HNSW_Grouping::write / read            (IndexIVF_HNSW_Grouping.cpp:397-483)
// Layout, every vector as [uint32 count][elements] (write_vector / read_vector, utils.h:52-66):
//   size_t d, nc, (nsubc);  ids[nc];  codes[nc];  norm_codes[nc];
//   (nn_centroid_idxs[nc];  subgroup_sizes[nc];  alphas);  centroid_norms;  (inter_centroid_dists[nc])
// The lists live in HBM (slot space); they are exported / loaded through ivfhnsw_b200_export_lists /
// ivfhnsw_b200_load_lists as flat CSR arrays and streamed to / from the file through one large buffer, so a
// 1B-vector index (22 GB, 3 x 2^20 vectors) is written in one pass without per-list allocations.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "handle.h"

namespace {

struct Writer {
    FILE *f = nullptr;
    std::vector<char> buf;
    size_t fill = 0;
    bool ok = true;
    explicit Writer(const char *path) : buf((size_t)64 << 20) { f = fopen(path, "wb"); ok = f != nullptr; }
    ~Writer() { if (f) fclose(f); }
    void flush()
    {
        if (fill && ok) ok = fwrite(buf.data(), 1, fill, f) == fill;
        fill = 0;
    }
    void put(const void *p, size_t n)
    {
        if (n >= buf.size() / 2) {
            flush();
            if (ok) ok = fwrite(p, 1, n, f) == n;
            return;
        }
        if (fill + n > buf.size()) flush();
        memcpy(buf.data() + fill, p, n);
        fill += n;
    }
    void vec(const void *p, size_t count, size_t elem)
    {
        const uint32_t c32 = (uint32_t)count;
        put(&c32, 4);
        if (count) put(p, count * elem);
    }
    bool close()
    {
        flush();
        if (f) { ok = (fclose(f) == 0) && ok; f = nullptr; }
        return ok;
    }
};

struct Reader {
    FILE *f = nullptr;
    bool ok = true;
    explicit Reader(const char *path)
    {
        f = fopen(path, "rb");
        ok = f != nullptr;
        if (f) setvbuf(f, nullptr, _IOFBF, (size_t)16 << 20);
    }
    ~Reader() { if (f) fclose(f); }
    void get(void *p, size_t n) { if (ok && n) ok = fread(p, 1, n, f) == n; }
    uint32_t count() { uint32_t c = 0; get(&c, 4); return ok ? c : 0u; }
};

}  // namespace

extern "C" {

int ivfhnsw_b200_write_index(ivfhnsw_b200_t *h, const char *path)
{
    if (!h || !path) return IVFHNSW_B200_EINVAL;
    if (int r = ivfhnsw_b200_finalize(h)) return r;
    const size_t nc = h->nc, M = h->M, ns = h->nsubc;
    size_t n = 0;
    ivfhnsw_b200_ntotal(h, &n);
    std::vector<int64_t> off(nc + 1);
    std::vector<uint32_t> ids(n ? n : 1), nn(ns ? nc * ns : 0), ss(ns ? nc * ns : 0);
    std::vector<uint8_t> codes((n ? n : 1) * M), ncodes(n ? n : 1);
    std::vector<float> alphas(ns ? nc : 0), icd(ns ? nc * ns : 0), cn(nc);
    if (int r = ivfhnsw_b200_export_lists(h, off.data(), ids.data(), codes.data(), ncodes.data(), ns ? nn.data() : nullptr,
                                          ns ? ss.data() : nullptr, ns ? alphas.data() : nullptr, ns ? icd.data() : nullptr))
        return r;
    if (int r = ivfhnsw_b200_get_centroid_norms(h, cn.data())) return r;
    Writer w(path);
    if (!w.ok) return h->fail(IVFHNSW_B200_EINVAL, "cannot open %s for writing", path);
    const size_t d = h->d;
    w.put(&d, sizeof(size_t));
    w.put(&nc, sizeof(size_t));
    if (ns) w.put(&ns, sizeof(size_t));
    for (size_t c = 0; c < nc; c++) w.vec(ids.data() + off[c], (size_t)(off[c + 1] - off[c]), 4);
    for (size_t c = 0; c < nc; c++) w.vec(codes.data() + (size_t)off[c] * M, (size_t)(off[c + 1] - off[c]) * M, 1);
    for (size_t c = 0; c < nc; c++) w.vec(ncodes.data() + off[c], (size_t)(off[c + 1] - off[c]), 1);
    if (ns) {
        for (size_t c = 0; c < nc; c++) w.vec(nn.data() + c * ns, ns, 4);
        // the reference leaves subgroup_sizes empty for groups that never received points
        for (size_t c = 0; c < nc; c++) w.vec(ss.data() + c * ns, off[c + 1] > off[c] ? ns : 0, 4);
        w.vec(alphas.data(), nc, 4);
    }
    w.vec(cn.data(), nc, 4);
    if (ns)
        for (size_t c = 0; c < nc; c++) w.vec(icd.data() + c * ns, ns, 4);
    if (!w.close()) return h->fail(IVFHNSW_B200_EINVAL, "write error on %s (disk full?)", path);
    return 0;
}

int ivfhnsw_b200_read_index(ivfhnsw_b200_t *h, const char *path)
{
    if (!h || !path) return IVFHNSW_B200_EINVAL;
    Reader r(path);
    if (!r.ok) return h->fail(IVFHNSW_B200_EINVAL, "cannot open index file %s", path);
    const size_t nc = h->nc, M = h->M, ns = h->nsubc;
    size_t fd = 0, fnc = 0, fns = 0;
    r.get(&fd, sizeof(size_t));
    r.get(&fnc, sizeof(size_t));
    if (ns) r.get(&fns, sizeof(size_t));
    if (!r.ok || fd != h->d || fnc != nc || (ns && fns != ns))
        return h->fail(IVFHNSW_B200_EINVAL, "index file %s is for d=%zu nc=%zu nsubc=%zu (this index: %zu, %zu, %zu)", path,
                       fd, fnc, fns, h->d, nc, ns);
    // pass 1 over the id section gives the list sizes; the other sections repeat them
    std::vector<int64_t> off(nc + 1, 0);
    std::vector<uint32_t> ids;
    {
        size_t cap = 1 << 20;
        ids.resize(cap);
        size_t n = 0;
        for (size_t c = 0; c < nc && r.ok; c++) {
            const uint32_t cnt = r.count();
            if (n + cnt > cap) {
                while (n + cnt > cap) cap *= 2;
                ids.resize(cap);
            }
            r.get(ids.data() + n, (size_t)cnt * 4);
            n += cnt;
            off[c + 1] = (int64_t)n;
        }
    }
    const size_t n = (size_t)off[nc];
    std::vector<uint8_t> codes((n ? n : 1) * M), ncodes(n ? n : 1);
    for (size_t c = 0; c < nc && r.ok; c++) {
        const uint32_t cnt = r.count();
        if ((size_t)cnt != (size_t)(off[c + 1] - off[c]) * M) { r.ok = false; break; }
        r.get(codes.data() + (size_t)off[c] * M, cnt);
    }
    for (size_t c = 0; c < nc && r.ok; c++) {
        const uint32_t cnt = r.count();
        if ((size_t)cnt != (size_t)(off[c + 1] - off[c])) { r.ok = false; break; }
        r.get(ncodes.data() + off[c], cnt);
    }
    std::vector<uint32_t> nn(ns ? nc * ns : 0, 0u), ss(ns ? nc * ns : 0, 0u);
    std::vector<float> alphas(ns ? nc : 0, 0.f), icd(ns ? nc * ns : 0, 0.f), cn;
    std::vector<uint32_t> tmp32;
    std::vector<float> tmpf;
    auto row32 = [&](uint32_t *dst) {
        const uint32_t cnt = r.count();
        tmp32.resize(cnt);
        r.get(tmp32.data(), (size_t)cnt * 4);
        if (r.ok) memcpy(dst, tmp32.data(), 4 * (cnt < ns ? cnt : ns));
    };
    if (ns) {
        for (size_t c = 0; c < nc && r.ok; c++) row32(nn.data() + c * ns);
        for (size_t c = 0; c < nc && r.ok; c++) row32(ss.data() + c * ns);
        const uint32_t cnt = r.count();
        tmpf.resize(cnt);
        r.get(tmpf.data(), (size_t)cnt * 4);
        if (r.ok) memcpy(alphas.data(), tmpf.data(), 4 * (cnt < nc ? cnt : nc));
    }
    {
        const uint32_t cnt = r.count();
        cn.resize(cnt);
        r.get(cn.data(), (size_t)cnt * 4);
    }
    if (ns)
        for (size_t c = 0; c < nc && r.ok; c++) {
            const uint32_t cnt = r.count();
            tmpf.resize(cnt);
            r.get(tmpf.data(), (size_t)cnt * 4);
            if (r.ok) memcpy(icd.data() + c * ns, tmpf.data(), 4 * (cnt < ns ? cnt : ns));
        }
    if (!r.ok) return h->fail(IVFHNSW_B200_EINVAL, "index file %s is truncated or inconsistent", path);
    if (ids.empty()) ids.resize(1);
    if (int rc = ivfhnsw_b200_load_lists(h, off.data(), ids.data(), codes.data(), ncodes.data(), ns ? nn.data() : nullptr,
                                         ns ? ss.data() : nullptr, ns ? alphas.data() : nullptr, ns ? icd.data() : nullptr))
        return rc;
    if (cn.size() == nc) return ivfhnsw_b200_set_centroid_norms(h, cn.data());
    return 0;
}

}  // extern "C"
