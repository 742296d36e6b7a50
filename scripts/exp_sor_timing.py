"""orbm_statistical_outlier_removal on map-sized clouds: time, fraction of tile pairs visited (coherent = voxel-index order)"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from orb_slam2_map_b200 import capi
import test_sor as T
for n, coherent in ((20000, True), (200000, True), (200000, False), (1000000, True)):
    p = T._cloud(n, 3, coherent=coherent)
    capi.statistical_outlier_removal(p[:1000], 50, 1.0)
    t0 = time.perf_counter(); kept, md, st = capi.statistical_outlier_removal(p, 50, 1.0); dt = time.perf_counter() - t0
    print(f"n={n} coherent={coherent}: {dt*1e3:.1f} ms, kept {len(kept)}, threshold {st[0]:.5f}, tile pairs visited {st[4]:.4f}", flush=True)
