"""ncu --csv launch list (gpu__time_duration.sum) -> 'kernel grid block us' lines."""
import csv, sys
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
hdr = rows[0]; ik, ig, ib, iv, iu = hdr.index("Kernel Name"), hdr.index("Grid Size"), hdr.index("Block Size"), hdr.index("Metric Value"), hdr.index("Metric Unit")
for r in rows[1:]:
    v = float(r[iv].replace(",", "")); v = v / 1e3 if r[iu] in ("ns", "nsecond") else v
    print(f"{r[ik].split('(')[0][:28]:28s} {r[ig]:>16s} {r[ib]:>14s} {v:8.2f} us")
