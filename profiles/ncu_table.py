#!/usr/bin/env python
"""Counter table of an .ncu-rep (one row per kernel launch): usage ncu_table.py report.ncu-rep [json_out]"""
import csv, io, json, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[0]
cols = [("gpu__time_duration.sum", "us", 1e-3 if rows[1][hdr.index("gpu__time_duration.sum")] in ("ns", "nsecond") else 1),
        ("dram__bytes_read.sum", "rdMB", None), ("dram__bytes_write.sum", "wrMB", None),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM%", 1),
        ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "L1%", 1),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2%", 1),
        ("lts__t_sector_hit_rate.pct", "L2hit%", 1),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps%", 1),
        ("sm__inst_issued.avg.pct_of_peak_sustained_active", "issue%", 1),
        ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "ALU%", 1),
        ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "LSU%", 1),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor%", 1),
        ("launch__registers_per_thread", "regs", 1),
        ("launch__occupancy_limit_shared_mem", "CTA/SM(smem)", 1), ("launch__occupancy_limit_registers", "CTA/SM(regs)", 1)]
unit = dict(zip(hdr, rows[1]))
def mb(v, u):
    v = float(v)
    return v * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1, "Gbyte": 1e3}.get(u, 1)
print("| kernel | " + " | ".join(c[1] for c in cols) + " |")
print("|---|" + "---:|" * len(cols))
traffic = {}
for r in rows[2:]:
    name = r[hdr.index("Kernel Name")].split("(")[0].replace("void ", "")
    vals = []
    for key, lab, sc in cols:
        if key not in hdr:
            vals.append("-"); continue
        v = r[hdr.index(key)]
        if sc is None:
            vals.append("%.1f" % mb(v, unit[key]))
        else:
            try:
                f = float(v)
                if key == "gpu__time_duration.sum":
                    f *= {"ns": 1e-3, "nsecond": 1e-3, "us": 1, "usecond": 1, "ms": 1e3, "msecond": 1e3}.get(unit[key], 1)
                vals.append("%.1f" % f)
            except ValueError:
                vals.append(v)
    print("| `%s` | " % name + " | ".join(vals) + " |")
    traffic[name.split("<")[0]] = int((mb(r[hdr.index("dram__bytes_read.sum")], unit["dram__bytes_read.sum"]) +
                                       mb(r[hdr.index("dram__bytes_write.sum")], unit["dram__bytes_write.sum"])) * 1e6)
if len(sys.argv) > 2:
    json.dump(traffic, open(sys.argv[2], "w"), indent=1)
