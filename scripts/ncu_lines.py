#!/usr/bin/env python
"""Attribute executed instructions / stall samples of one kernel in an .ncu-rep to CUDA source lines.

usage: ncu_lines.py <report.ncu-rep> <kernel-regex> [liborbx.so] [--top N] [--id K]

ncu's `--page source --csv` is per SASS instruction; the line table comes from `nvdisasm -g` on the
cubin inside the .so that was profiled (built with -lineinfo).  Both list the kernel's instructions in
address order, so they are joined by position.  Runs on the CPU box (no GPU needed).
"""
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile


def sass_lines(so, kernel_re):
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, check=True, stdout=subprocess.DEVNULL)
    out = []
    for f in sorted(os.listdir(tmp)):
        if not f.endswith(".cubin"):
            continue
        txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, f)], capture_output=True, text=True).stdout
        cur, line, active = None, 0, False
        for ln in txt.splitlines():
            m = re.match(r"\s*//-+ \.text\.(\S+)", ln)
            if m:
                active = re.search(kernel_re, m.group(1)) is not None
                if active and out:
                    return out          # first matching function only
                continue
            if not active:
                continue
            m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
            if m:
                line = int(m.group(2)); cur = m.group(1)
                continue
            m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
            if m:
                out.append((int(m.group(1), 16), line, cur, m.group(2).strip()))
        if out:
            return out
    return out


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    rep, kre = args[0], args[1]
    so = args[2] if len(args) > 2 else "orb_slam2_map_b200/liborbx.so"
    top = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 40
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "-k", "regex:" + kre],
                         capture_output=True, text=True).stdout
    # the report may hold several launches of the kernel: split on the "Kernel Name" header rows
    blocks, cur = [], None
    for r in csv.reader(io.StringIO(raw)):
        if r and r[0] == "Kernel Name":
            cur = []; blocks.append(cur); continue
        if cur is not None:
            cur.append(r)
    kid = int(sys.argv[sys.argv.index("--id") + 1]) if "--id" in sys.argv else 0
    rows = blocks[kid]
    hdr = rows[0]
    ci, si = hdr.index("Instructions Executed"), hdr.index("# Samples")
    ti = hdr.index("Thread Instructions Executed")
    inst = [(r[1], int(r[ci]), int(r[si]), int(r[ti])) for r in rows[1:] if len(r) > ci and r[ci].isdigit()]
    lines = sass_lines(so, kre)
    if len(lines) != len(inst):
        print(f"warning: {len(lines)} SASS instructions in {so} vs {len(inst)} in the report (different build?)")
    n = min(len(lines), len(inst))
    per = collections.defaultdict(lambda: [0, 0, 0])
    tot_i = tot_s = 0
    for k in range(n):
        _, ln, fn, _ = lines[k]
        ln = (fn, ln)                       # inlined header code keeps its own file
        per[ln][0] += inst[k][1]; per[ln][1] += inst[k][2]; per[ln][2] += inst[k][3]
        tot_i += inst[k][1]; tot_s += inst[k][2]
    src = {}
    for f in {l[2] for l in lines if l[2]}:
        if os.path.exists(f):
            for i, l in enumerate(open(f)):
                src[(f, i + 1)] = l.rstrip()
    main_file = lines[0][2] if lines else None
    print(f"kernel {kre}: {tot_i} warp instructions, {tot_s} stall samples, {n} SASS instructions")
    print(f"{'inst%':>6} {'smp%':>6} {'lanes':>5}  line  source")
    items = sorted(per.items()) if "--by-line" in sys.argv else sorted(per.items(), key=lambda kv: -kv[1][0])[:top]
    for ln, (i, s, t) in items:
        tag = "" if ln[0] == main_file else "[" + os.path.basename(ln[0] or "?") + "] "
        print(f"{100 * i / max(tot_i, 1):6.2f} {100 * s / max(tot_s, 1):6.2f} {t / max(i, 1):5.1f}  {ln[1]:4d}  {tag}{src.get(ln, '')[:110]}")


if __name__ == "__main__":
    main()
