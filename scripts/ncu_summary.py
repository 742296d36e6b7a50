#!/usr/bin/env python
"""Summarise an .ncu-rep (read on the CPU box): per kernel time, DRAM bytes, pipe/issue utilisation, top stalls."""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
def col(n): return hdr.index(n) if n in hdr else None
base = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "smsp__inst_executed.sum",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "smsp__thread_inst_executed_per_inst_executed.ratio"]
stalls = [h for h in hdr if h.startswith("smsp__average_warp") and "issue_stalled" in h and h.endswith("_per_warp_active.pct")]
if not stalls:
    stalls = [h for h in hdr if "warp_issue_stalled" in h and h.endswith(".pct")]
for r in rows[2:]:
    print("==", r[col("Kernel Name")][:60])
    for b in base:
        c = col(b)
        if c is not None: print(f"   {b:68s} {r[c]:>16s} {units[c]}")
    st = []
    for s in stalls:
        try: st.append((float(r[col(s)]), s))
        except ValueError: pass
    for v, s in sorted(st, reverse=True)[:6]:
        print(f"   stall {s.split('issue_stalled_')[1][:40]:42s} {v:8.2f}")
