"""TEST INFRASTRUCTURE -- generate tests/golden/*.npz from the COMPILED REFERENCE.

Run in the build container (where /root/reference exists):
    make -C oracle ref && python oracle/make_golden.py
The reference ships no golden vectors of its own, so these files
are the pin: every array below was produced by ivfhnsw::IndexIVF_HNSW /
IndexIVF_HNSW_Grouping themselves (through oracle/ref_shim.cpp), following the
call order of tests/test_ivfhnsw_sift1b.cpp:44-208 and
tests/test_ivfhnsw_grouping_sift1b.cpp.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import oracle as O  # noqa: E402
import synth_np  # noqa: E402

GOLD = os.path.join(os.path.dirname(HERE), "tests", "golden")

#        name            kind    d    nc   M  nsubc opq   nb    nt    nq nprobe k   efS
CASES = [
    ("adc_d32",         "sift",  32,  64,  8,  0, False, 3000, 2500, 40,  8, 10, 64),
    ("adc_opq_d32",     "sift",  32,  64,  8,  0, True,  3000, 2500, 40,  8, 10, 64),
    ("adc_sift_m8",     "sift", 128, 128,  8,  0, False, 4000, 3000, 40, 16, 10, 128),
    ("adc_sift_m16",    "sift", 128, 128, 16,  0, False, 4000, 3000, 40, 16, 100, 128),
    ("enn_sift_m64",    "sift", 128,  64, 64,  0, False, 2000, 2500, 32,  1,  1, 64),
    ("grp_d32",         "sift",  32,  64,  8,  8, False, 3000, 2500, 40,  8, 10, 64),
    ("grp_deep_opq",    "deep",  96,  96, 16, 16, True,  4000, 3000, 40, 12, 10, 96),
    ("grp_sift_m16",    "sift", 128, 128, 16, 32, False, 5000, 3000, 40, 16, 10, 128),
    # long lists and sub-groups (~750 / ~94 codes): the scan tiles runs of kept sub-groups instead of whole lists
    ("grp_long_m16",    "sift",  64,  32, 16,  8, False, 24000, 3000, 40, 8, 10, 32),
]
NSUB = 400  # base vectors kept in the fixture for indexing parity


def run_case(name, kind, d, nc, M, nsubc, opq, nb, nt, nq, nprobe, k, efs, seed):
    base, queries, learn, centroids = synth_np.dataset(kind, nb, nq, nt, nc, d, seed)
    if name.startswith("enn"):
        queries = base[:nq].copy()  # tests/test_ivfhnsw_sift1b_enn.cpp:27-47
    r = O.Ref(d, nc, M, nsubc)
    r.build_quantizer(centroids, M=16, efConstruction=200)
    r.set_do_opq(opq)
    r.set_params(nprobe, 10 ** 9, 220, False)
    r.train_pq(learn)
    ids = np.arange(nb, dtype=np.uint32)
    idx = r.assign(base)                      # HNSW assign, efSearch = 220 (tests/...sift1b.cpp:108)
    out = dict(d=d, nc=nc, M=M, nsubc=nsubc, opq_on=int(opq), nprobe=nprobe, k=k, seed=seed,
               kind=kind, centroids=centroids, X_sub=base[:NSUB], assign_ref=idx[:NSUB], assign_all=idx)
    if nsubc == 0:
        r.add_batch(base, ids, idx)
    else:
        for g in range(nc):
            m = idx == g
            r.add_group(g, base[m], ids[m])
    r.compute_centroid_norms()
    if nsubc:
        r.compute_inter_centroid_dists()
    if opq:
        r.rotate_quantizer()
    ix = r.export()
    out["centroids_search"] = ix["centroids"]  # rotated in place when opq (IndexIVF_HNSW.cpp:789-800)
    for key in ("pqc", "norm_table", "centroid_norms", "list_off", "ids", "codes", "norm_codes"):
        out[key] = ix[key]
    if opq:
        out["opq"] = ix["opq"]
    if nsubc:
        for key in ("nn_idx", "sub_sizes", "alphas", "icd"):
            out[key] = ix[key]
    out["Q"] = queries
    avg = max(nb // nc, 1)
    runs = []
    for max_codes in (10 ** 9, avg * max(nprobe // 3, 1)):
        for pruning in ((False, True) if nsubc else (False,)):
            r.set_params(nprobe, max_codes, efs, pruning)
            D, L, cd, ci = r.search(queries, k)
            tag = "mc%d_pr%d" % (0 if max_codes == 10 ** 9 else max_codes, int(pruning))
            out["D_" + tag], out["L_" + tag] = D, L
            out["cd_" + tag], out["ci_" + tag] = cd, ci
            runs.append((tag, max_codes, int(pruning)))
            if nsubc == 0:  # search2 must agree with search on the same candidates
                D2, L2 = r.search2(queries, k, cd, ci)
                assert np.array_equal(L, L2) and np.array_equal(D, D2)
    out["runs"] = np.array([t for t, _, _ in runs])
    out["runs_max_codes"] = np.array([m for _, m, _ in runs], np.int64)
    out["runs_pruning"] = np.array([p for _, _, p in runs], np.int64)
    r.close()
    np.savez_compressed(os.path.join(GOLD, name + ".npz"), **out)
    print("wrote", name, "N=%d" % ix["ids"].size, "runs", [t for t, _, _ in runs])


if __name__ == "__main__":
    O.build()
    os.makedirs(GOLD, exist_ok=True)
    only = sys.argv[1:]
    for i, case in enumerate(CASES):
        if only and case[0] not in only:
            continue
        run_case(*case, seed=1234 + i)
