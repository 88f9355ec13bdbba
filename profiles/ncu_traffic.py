#!/usr/bin/env python
"""DRAM traffic per launch of every kernel of an `ncu --set full` capture -> profiles/r02_ncu_traffic.json, tagged with
the sha1 of the kernel sources it was captured from (bench.py refuses a file whose hash differs from the tree's).

    python profiles/ncu_traffic.py gpurun_out/c4s_full.ncu-rep c4s [gpurun_out/gemm.ncu-rep c3 ...]

traffic = dram__bytes_read.sum + dram__bytes_write.sum, averaged over the launches of a kernel in the capture."""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def traffic_of(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, unit = rows[0], dict(zip(rows[0], rows[1]))
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    acc = {}
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")].split("(")[0].replace("void ", "").split("<")[0].split("::")[-1]
        b = sum(float(r[hdr.index(k)]) * scale.get(unit[k], 1.0) for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
        acc.setdefault(name, []).append(b)
    return {k: int(sum(v) / len(v)) for k, v in acc.items()}


def main():
    import bench
    path = os.path.join(ROOT, "profiles", "r02_ncu_traffic.json")
    try:
        cur = json.load(open(path))
    except Exception:
        cur = {}
    h = bench.csrc_hash()
    if cur.get("csrc_sha1") != h:
        cur = {"csrc_sha1": h, "note": "bytes per launch, dram__bytes_read.sum + dram__bytes_write.sum, ncu --set full"}
    args = sys.argv[1:]
    for rep, wl in zip(args[0::2], args[1::2]):
        cur.setdefault(wl, {}).update(traffic_of(rep))
    json.dump(cur, open(path, "w"), indent=1)
    print(json.dumps(cur, indent=1))


if __name__ == "__main__":
    main()
