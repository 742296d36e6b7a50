"""Repeat ONE stateless matcher call (argv[1]: init | bow | sim3) a few times: the target of an ncu capture."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import oracle_lib as O
from orb_slam2_map_b200 import capi, synth

which = sys.argv[1]
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 6
if which == "init":
    ex = capi.Extractor(max_width=640, max_height=480, chunk_frames=1, result_slots=1)
    fr = synth.frame_sequence(2, 640, 480, seed=3)
    ka, da = ex.extract(fr[0]); kb, db = ex.extract(fr[1])
    f1, f2 = capi.FrameData(ka, da, 640, 480), capi.FrameData(kb, db, 640, 480)
    prev = np.stack([ka["x"], ka["y"]], axis=1)
    for _ in range(reps):
        capi.search_for_initialization(f1, f2, prev, 100, 0.9, True)
elif which == "bow":
    from test_oracle_bow import make_pair
    voc = O.make_vocabulary(k=10, depth=4, seed=3)
    A, B = make_pair(voc, n=1000, seed=1, levelsup=2)
    for _ in range(reps):
        capi.search_by_bow(0, [(A, B)], nnratio=0.7, check_ori=True)
