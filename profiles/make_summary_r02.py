#!/usr/bin/env python
"""Tables of r02_ncu_summary.md: usage make_summary_r02.py launches.csv full.ncu-rep  (text around them is by hand)."""
import csv
import os
import subprocess
import sys

here = os.path.dirname(os.path.abspath(__file__))


def launch_table(f):
    rows = [r for r in csv.reader(open(f)) if len(r) > 10]
    hdr = rows[0]
    ki, vi, gi, bi = (hdr.index(c) for c in ("Kernel Name", "Metric Value", "Grid Size", "Block Size"))
    body = rows[1:]
    # the LAST search step of the capture: from the last rotate/split launch that starts a step to the end of its scan
    names = [r[ki] for r in body]
    last_scan = max(i for i, n in enumerate(names) if "scan_kernel" in n)
    start = max(i for i, n in enumerate(names[:last_scan]) if "rotate_kernel" in n or "split_rows_kernel" in n)
    # a step begins with rotate (OPQ) if present, else split_rows
    while start > 0 and ("rotate_kernel" in names[start - 1] or "FillFunctor" in names[start - 1]):
        start -= 1
    sel = [r for r in body[start:last_scan + 1] if "FillFunctor" not in r[ki] and "elementwise" not in r[ki]]
    tot = sum(float(r[vi]) for r in sel)
    out = ["| kernel | grid x block | ncu us | share |", "|---|---|---:|---:|"]
    for r in sel:
        name = r[ki].split("(")[0].replace("void ", "").replace("ivfb200::", "").replace("<unnamed>::", "")
        out.append("| `%s` | %s x %s | %.1f | %.1f %% |" % (name, r[gi].strip("()").split(",")[0],
                                                           r[bi].strip("()").split(",")[0], float(r[vi]) / 1e3,
                                                           100 * float(r[vi]) / tot))
    out.append("| **step** | | **%.0f** | |" % (tot / 1e3))
    return "\n".join(out)


print("## launch list (one timed step)\n")
print(launch_table(sys.argv[1]))
print("\n## counters (ncu --set full, one launch each)\n")
print(subprocess.run([sys.executable, os.path.join(here, "ncu_table.py"), sys.argv[2]], capture_output=True, text=True).stdout)
