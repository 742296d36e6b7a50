#!/bin/bash
# wall time of the two default bench arms (the driver runs them with no flags)
python - <<'PY'
import subprocess, time
for cmd in (["python", "bench.py"], ["python", "bench.py", "--impl", "reference"]):
    t = time.time(); r = subprocess.run(cmd, capture_output=True, text=True); print(" ".join(cmd), "rc", r.returncode, "wall %.1f s" % (time.time() - t), "stdout lines", len(r.stdout.strip().splitlines()))
    print(r.stdout.strip()[:700])
PY
