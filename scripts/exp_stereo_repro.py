"""repro of soak trial (seed 7, t 336): ComputeStereoMatches 526x300 nf 2000"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import oracle_lib as O
from orb_slam2_map_b200 import capi, synth
w, h, nf, seed, t = 526, 300, 2000, 7, 336
left, right = synth.stereo_pair(w, h, seed=seed + t)
el, er = O.Extractor(nf), O.Extractor(nf)
kl, dl = el.extract(left); kr, dr = er.extract(right)
pl = [el.level(l) for l in range(8)]; pr = [er.level(l) for l in range(8)]
tb = el.tables()
bf, fx = 386.1448, 718.856
u_o, z_o, c_o = O.compute_stereo_matches(kl, dl, kr, dr, pl, pr, tb["scale"], tb["inv_scale"], bf, fx)
for rep in range(3):
    u_g, z_g, c_g = capi.compute_stereo_matches(kl, dl, kr, dr, pl, pr, tb["scale"], tb["inv_scale"], bf, fx)
    bad = np.nonzero((u_g != u_o) | (z_g != z_o))[0]
    print("rep", rep, "n", len(kl), len(kr), "compares", c_g, c_o, "differ", len(bad), bad[:10], u_g[bad[:5]], u_o[bad[:5]], "matched", (u_o >= 0).sum(), (u_g >= 0).sum())
