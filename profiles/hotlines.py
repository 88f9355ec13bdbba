#!/usr/bin/env python
"""Per-source-line hot spots from an .ncu-rep (needs -lineinfo + --import-source on).
usage: hotlines.py report.ncu-rep kernel_regex [top_n]"""
import csv, subprocess, sys, io

rep, kre = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv",
                      "--kernel-name", "regex:" + kre], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
fname, hdr, lines = "", None, []
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        fname = r[1].split("/")[-1]
    elif len(r) > 5 and r[0] == "Line No":
        hdr = r
        ii = hdr.index("Instructions Executed"); wi = hdr.index("Warp Stall Sampling (All Samples)")
    elif hdr and len(r) > 5 and r[0] not in ("", "Line No"):
        try:
            lines.append((fname, int(r[0]), r[1].strip(), int(r[ii] or 0), int(r[wi] or 0)))
        except ValueError:
            pass
ti = sum(l[3] for l in lines) or 1
tw = sum(l[4] for l in lines) or 1
print(f"# {kre}: warp-instructions {ti}, stall samples {tw}")
print("# inst%  stall%  file:line  source")
by = 3 if len(sys.argv) > 4 and sys.argv[4] == "inst" else 4
for l in sorted(lines, key=lambda x: -x[by])[:top]:
    print(f"{l[3]/ti*100:6.2f} {l[4]/tw*100:6.2f}  {l[0]}:{l[1]}  {l[2][:100]}")
