"""One GPU standing in for one rank of an N-way list-sharded deployment: the handle scans 1/N of the lists for ALL queries
(set_shard), which is the per-rank shape of the scan in the sharded layout.  Prints the device time of search_dev per
shard fraction; under ncu (--nvtx-include "probe/") the scan kernel of the 1/8 case is what gets captured."""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402


def main():
    wl = sys.argv[1] if len(sys.argv) > 1 else "c4s"
    w = dict(bench.WORKLOADS[wl])
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    log = lambda m: print("[probe] " + m, file=sys.stderr, flush=True)
    idx, queries, gt0, _ = bench.build_index(w, torch, dev, log)
    nq, k = w["nq"], w["k"]
    D = torch.empty(nq, k, device=dev)
    L = torch.empty(nq, k, dtype=torch.int64, device=dev)
    for parts in (1, 2, 4, 8):
        idx.set_shard(0, w["nc"] // parts)
        for _ in range(3):
            idx.search_dev(k, queries, D, L)
        torch.cuda.synchronize()
        idx.reset_stats()
        idx.set_timing(True)
        for _ in range(3):
            idx.search_dev(k, queries, D, L)
        torch.cuda.synchronize()
        st = idx.stats()
        idx.set_timing(False)
        print("shard 1/%d: scan %.3f ms (lut %.3f) per %d queries, %.0f scanned codes/query" %
              (parts, st.scan_ms / 3, st.lut_ms / 3, nq, st.scanned_codes / 3 / nq), flush=True)
    idx.set_shard(0, w["nc"] // 8)
    torch.cuda.nvtx.range_push("probe")
    idx.search_dev(k, queries, D, L)
    torch.cuda.synchronize()
    torch.cuda.nvtx.range_pop()


if __name__ == "__main__":
    main()
