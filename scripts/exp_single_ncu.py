"""Driver for `ncu` on the single-frame path: N calls of orbx_extract on one handle (graph replay after the first)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from orb_slam2_map_b200 import capi, synth
frames = synth.frame_sequence(4, 640, 480, seed=0)
ex = capi.Extractor(max_width=640, max_height=480, chunk_frames=1, result_slots=1, device=0)
for i in range(int(sys.argv[1]) if len(sys.argv) > 1 else 6):
    k, d = ex.extract(frames[i % 4])
print(len(k))
