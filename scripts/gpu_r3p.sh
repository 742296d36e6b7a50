#!/bin/bash
mkdir -p gpurun_out
python bench.py > gpurun_out/r02z_bench.json 2> gpurun_out/r02z_bench.err; echo "bench rc=$?"
python bench.py --impl reference > gpurun_out/r02z_bench_ref.json 2> gpurun_out/r02z_bench_ref.err; echo "ref rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02z_bench.json").read().strip().splitlines()[-1]); c=d["cpu_baseline"]
print(round(d["value"]), round(d["e2e"]["value"]), "traffic", d["roofline"]["traffic"] is not None, "cpu", round(c["value"]), round(c.get("value_o3_native",0)), c["sample"][:90], c["cores"])
r=json.loads(open("gpurun_out/r02z_bench_ref.json").read().strip().splitlines()[-1]); print("ref", round(r["value"]), r["config"]["frames_per_step"])
PY
