#!/usr/bin/env python
"""Regenerates the tables of r01_ncu_summary.md from the capture directory written by scratch/final_capture.sh
(usage: make_summary.py gpurun_out/final > /tmp/tables.md).  Text around the tables is edited by hand."""
import csv
import json
import os
import subprocess
import sys

src = sys.argv[1]
here = os.path.dirname(os.path.abspath(__file__))


def launch_table(f):
    rows = [r for r in csv.reader(open(f)) if len(r) > 10]
    hdr = rows[0]
    ki, vi, gi, bi = (hdr.index(c) for c in ("Kernel Name", "Metric Value", "Grid Size", "Block Size"))
    n = len(rows[1:]) // 2          # two timed steps were captured: keep the second
    sel = [r for r in rows[1 + n:] if "FillFunctor" not in r[ki]]
    tot = sum(float(r[vi]) for r in sel)
    out = ["| kernel | grid x block | ncu us | share |", "|---|---|---:|---:|"]
    for r in sel:
        name = r[ki].split("(")[0].replace("void ", "").replace("ivfb200::", "").replace("<unnamed>::", "")
        out.append("| `%s` | %s x %s | %.1f | %.1f %% |" % (name, r[gi].strip("()").split(",")[0],
                                                           r[bi].strip("()").split(",")[0], float(r[vi]) / 1e3,
                                                           100 * float(r[vi]) / tot))
    out.append("| **step** | | **%.0f** | |" % (tot / 1e3))
    return "\n".join(out)


traffic = {}
for w in ("c2", "c4s"):
    print("## launch list %s\n" % w)
    print(launch_table(os.path.join(src, "launches_%s.csv" % w)))
    print("\n## counters %s\n" % w)
    tj = "/tmp/traffic_%s.json" % w
    print(subprocess.run([sys.executable, os.path.join(here, "ncu_table.py"), os.path.join(src, "full_%s.ncu-rep" % w), tj],
                         capture_output=True, text=True).stdout)
    traffic[w] = json.load(open(tj))
traffic["_note"] = ("dram__bytes_read.sum + dram__bytes_write.sum per launch from the round-1 `ncu --set full` captures "
                    "(one 10k-query step; profiles/r01_ncu_summary.md)")
json.dump(traffic, open(os.path.join(here, "r01_ncu_traffic.json"), "w"), indent=1)
for w in ("c2", "c4s", "c1", "c3", "c5s"):
    p = os.path.join(src, "bench_%s.json" % w)
    if os.path.exists(p):
        l = json.load(open(p))
        print("bench %s: value %.4g %s, %.3f ms/step, e2e %.4g, recall %s, roofline %s frac %.3f, cpu %s" % (
            w, l["value"], l["unit"], l["ms_per_step"], l["e2e"]["value"], l.get("recall_at_10"), l["roofline"]["kernel"],
            l["roofline"]["frac"], (l.get("cpu_baseline") or {}).get("value")))
