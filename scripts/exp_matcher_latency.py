"""Latency of the stateless per-frame matcher calls (host structs in, host results out) next to the CPU oracle:
SearchByProjection (900 projected map points against one 640x480 frame) and SearchForInitialization (one pair)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import oracle_lib as O
from orb_slam2_map_b200 import capi, synth
import test_match2_gpu as T

def med(fn, n=60, warm=5):
    for _ in range(warm): fn()
    ts = []
    for _ in range(n):
        t0 = time.perf_counter(); fn(); ts.append(time.perf_counter() - t0)
    return float(np.median(ts)) * 1e6

k, d, q, qd, ur, occ, blk, sf = T._frame_with_queries(1)
fo = O.FrameData(k, d, 640, 480, u_right=ur, scale_factors=sf)
fg = capi.FrameData(k, d, 640, 480, u_right=ur, scale_factors=sf)
out = {}
for name, variant, th, ratio, ori in (("localmap", capi.PROJ_LOCALMAP, 100, 0.8, False), ("lastframe", capi.PROJ_LASTFRAME, 100, 0.9, True)):
    g = med(lambda: capi.search_by_projection(variant, fg, q, qd, blk, occ, th, ratio, ori))
    c = med(lambda: O.search_by_projection(variant, fo, q, qd, blk, occ, th, ratio, ori), n=20)
    out[f"search_by_projection_{name}"] = {"gpu_us": round(g, 1), "cpu_oracle_us": round(c, 1), "queries": len(q), "keypoints": len(k)}
frames = synth.frame_sequence(2, 640, 480, seed=3)
orc = O.Extractor()
ka, da = orc.extract(frames[0]); kb, db = orc.extract(frames[1])
f1o, f2o = O.FrameData(ka, da, 640, 480), O.FrameData(kb, db, 640, 480)
f1g, f2g = capi.FrameData(ka, da, 640, 480), capi.FrameData(kb, db, 640, 480)
prev = np.stack([ka["x"], ka["y"]], axis=1)
g = med(lambda: capi.search_for_initialization(f1g, f2g, prev, 100, 0.9, True))
c = med(lambda: O.search_for_initialization(f1o, f2o, prev, 100, 0.9, True), n=20)
out["search_for_initialization"] = {"gpu_us": round(g, 1), "cpu_oracle_us": round(c, 1)}
import json
print(json.dumps({"scratch_pool": not os.environ.get("ORBX_NO_SCRATCH_POOL"), **out}))
