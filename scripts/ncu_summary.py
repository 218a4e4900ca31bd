#!/usr/bin/env python
"""One line per captured launch of an .ncu-rep: time, DRAM bytes, throughputs, lanes, occupancy.
  python scripts/ncu_summary.py <report.ncu-rep>"""
import csv, io, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units, data = rows[0], rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
def g(r, k, scale=1.0):
    try: return float(r[ix[k]]) * scale
    except Exception: return float("nan")
def to_b(r, k):
    u = units[ix[k]]; v = g(r, k)
    return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}.get(u, 1)
def to_ms(r, k):
    u = units[ix[k]]; v = g(r, k)
    return v * {"ms": 1, "us": 1e-3, "ns": 1e-6, "s": 1e3}.get(u, 1)
print("kernel | ms | dram rd MB | dram wr MB | dram GB/s | sm% | issue% | lanes | warps_active% | regs | L1 hit% | L2 hit%")
for r in data:
    ms = to_ms(r, "gpu__time_duration.sum")
    rd, wr = to_b(r, "dram__bytes_read.sum"), to_b(r, "dram__bytes_write.sum")
    name = r[ix["Kernel Name"]].split("(")[0]
    print(f"{name} | {ms:.3f} | {rd/1e6:.1f} | {wr/1e6:.1f} | {(rd+wr)/ms/1e6:.0f} | "
          f"{g(r,'sm__throughput.avg.pct_of_peak_sustained_elapsed'):.1f} | {g(r,'smsp__issue_active.avg.pct_of_peak_sustained_active'):.1f} | "
          f"{g(r,'smsp__thread_inst_executed_per_inst_executed.ratio'):.1f} | {g(r,'sm__warps_active.avg.pct_of_peak_sustained_active'):.1f} | "
          f"{g(r,'launch__registers_per_thread'):.0f} | {g(r,'l1tex__t_sector_hit_rate.pct'):.1f} | {g(r,'lts__t_sector_hit_rate.pct'):.1f}")
