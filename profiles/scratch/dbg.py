import sys; sys.path.insert(0,'tests'); sys.path.insert(0,'.')
import numpy as np
from conftest import load_golden, load_pkg
from test_search_gpu import make_index
pkg = load_pkg()
for name in ["adc_d32", "enn_sift_m64", "adc_sift_m16"]:
    g, ix, runs = load_golden(name)
    idx = make_index(pkg, ix); idx.load_index_dict(ix); idx.nprobe = int(g["nprobe"])
    idx.set_option("debug", 2)
    for split in (1, 3):
        idx.set_option("scan_split", split)
        tag, mc, pr = runs[0]
        idx.max_codes, idx.do_pruning = mc, pr
        try:
            D, L = idx.search2(int(g["k"]), g["Q"], g["cd_" + tag], g["ci_" + tag])
            print(name, split, "ok", int(g["k"]), ix["M"], ix["d"])
        except Exception as e:
            print(name, split, "FAIL", str(e)[:300]); sys.exit(0)
