#!/usr/bin/env python
"""Condense an ncu launch list (--metrics gpu__time_duration.sum --csv) into per-kernel totals and shares."""
import collections, csv, re, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 14 and r[0].isdigit()]
tot = collections.OrderedDict()
for r in rows:
    name = re.sub(r"\(.*", "", r[4]).replace("void ", "").strip()[:60]
    t = tot.setdefault(name, [0, 0.0]); t[0] += 1; t[1] += float(r[-1].replace(",", "")) / 1e3
ours = re.compile(r"pyr_|blur7|fast_|octree|describe|grid_build|search_|stereo|knn2|repitch")
s = sum(v[1] for k, v in tot.items() if ours.search(k))
print(f"# {sys.argv[1]}: {len(rows)} launches; liborbx kernels total {s/1e3:.3f} ms (cold-cache, serialised)")
print(f"{'kernel':60s} {'launches':>8s} {'us total':>12s} {'us/launch':>10s} {'share':>7s}")
for k, v in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    sh = f"{100*v[1]/s:6.1f}%" if ours.search(k) else "   n/a"
    print(f"{k:60s} {v[0]:8d} {v[1]:12.1f} {v[1]/v[0]:10.1f} {sh}")
