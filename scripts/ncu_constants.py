#!/usr/bin/env python
"""profiles/ncu_constants.json from ONE `ncu --set full` capture of the extractor kernels of a bench chunk.

usage: ncu_constants.py <report.ncu-rep> <frames in the captured chunk> [label]

Per stage (resize = sum over its 7 launches): dram__bytes_read.sum + dram__bytes_write.sum and smsp__inst_executed.sum,
divided by the frames of the chunk; plus the sha256 of the kernel sources the report was captured on, so that bench.py can
tell a stale capture from a current one (it then reports traffic / issue as null)."""
import csv, hashlib, io, json, os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
STAGE = {"pyr_resize": "resize", "blur7": "blur", "fast_cells": "fast", "octree": "octree", "describe": "describe",
         "grid_build": "grid", "search_init": "search_init"}
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "inst": 1.0, "": 1.0}


def csrc_digest():
    d = os.path.join(ROOT, "orb_slam2_map_b200", "csrc")
    h = hashlib.sha256()
    for name in sorted(os.listdir(d)):
        if name.endswith((".cu", ".cuh", ".h")):
            with open(os.path.join(d, name), "rb") as f:
                h.update(name.encode()); h.update(f.read())
    return h.hexdigest()


def main():
    rep, frames = sys.argv[1], int(sys.argv[2])
    label = sys.argv[3] if len(sys.argv) > 3 else os.path.basename(rep)
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    col = {n: hdr.index(n) for n in ("Kernel Name", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
                                     "gpu__time_duration.sum")}
    val = lambda r, n: float(r[col[n]].replace(",", "")) * UNIT.get(units[col[n]], 1.0)
    out = {}
    for r in rows[2:]:
        stage = next((s for k, s in STAGE.items() if k in r[col["Kernel Name"]]), None)
        if stage is None:
            continue
        e = out.setdefault(stage, {"dram_bytes_per_frame": 0.0, "warp_inst_per_frame": 0.0, "launches": 0})
        e["dram_bytes_per_frame"] += (val(r, "dram__bytes_read.sum") + val(r, "dram__bytes_write.sum")) / frames
        e["warp_inst_per_frame"] += val(r, "smsp__inst_executed.sum") / frames
        e["launches"] += 1
    doc = {"source": f"profiles/{label}", "chunk_frames": frames, "image": "640x480, nfeatures 1000", "csrc_sha256": csrc_digest(), "kernels": out}
    with open(os.path.join(ROOT, "profiles", "ncu_constants.json"), "w") as f:
        json.dump(doc, f, indent=1)
    for k, v in out.items():
        print(f"{k:12s} {v['launches']} launches  dram {v['dram_bytes_per_frame'] / 1e6:7.3f} MB/frame  {v['warp_inst_per_frame'] / 1e6:6.3f} M warp-inst/frame")


if __name__ == "__main__":
    main()
