#!/usr/bin/env python
"""Randomised parity soak of the stateless matcher calls (orbm_* through the C ABI) against the CPU oracle: SearchByProjection x4
(latency form up to 4096 keypoints, grid + warp replay beyond), SearchForInitialization, SearchByBoW x2, Fuse x3,
SearchForTriangulation, ComputeDistinctiveDescriptors, on random sizes around the kernels' internal limits (0, 1, 31..33,
1023..1025, 2047..2049, 4095..4097, a few thousand).  Exits non-zero at the first mismatch.

usage: soak_matchers.py [--trials N] [--seed S] [--budget-s SECONDS]"""
import argparse, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import oracle_lib as O
from orb_slam2_map_b200 import capi
import test_match2_gpu as T2
from test_oracle_kfmatch import make_fuse_case, make_tri_pair, SCALE
from test_oracle_bow import make_pair
from test_kfmatch_gpu import _to_capi

SIZES = [0, 1, 2, 31, 32, 33, 100, 300, 700, 1000, 1023, 1024, 1025, 1500, 2047, 2048, 2049, 3000, 4095, 4096, 4097, 5000]


def pick(rng, lo=0, hi=5000, small=False):
    c = [s for s in SIZES if lo <= s <= hi]
    if small:
        c = [s for s in c if s <= 2049]
    return int(rng.choice(c))


def trial(t, seed, voc):
    rng = np.random.default_rng([seed, t])
    kind = ["proj", "proj", "init", "bow", "fuse", "tri", "distinct"][t % 7]
    if kind == "proj":
        n, nq = pick(rng, 1), pick(rng, 0, 5000)
        variant = int(rng.integers(0, 4))
        radius = float(rng.choice([5.0, 25.0, 70.0, 1000.0]))
        k, d = T2._random_frame(n, seed * 7919 + t)
        if n >= 3 and rng.integers(0, 2):                      # duplicated descriptors: ties across cells
            d[::3] = d[0]
        q, qd, occ, blk = T2._queries_around(k, d, nq, seed + t, radius)
        if variant >= 2:
            q["level_lo"] = rng.integers(0, 3, nq); q["level_hi"] = q["level_lo"] + 1
        else:
            lv = rng.integers(0, 8, nq)
            q["min_level"] = np.where(rng.random(nq) < 0.5, lv - 1, -1); q["max_level"] = np.where(rng.random(nq) < 0.7, lv + 1, -1)
        ur = np.where(rng.random(n) < 0.5, k["x"] - 10, -1.0).astype(np.float32) if rng.integers(0, 2) else None
        if ur is not None:
            q["u_right"] = q["u"] - 10; q["max_right_err"] = np.where(rng.random(nq) < 0.5, 4.0, -1.0)
        th, ratio, ori = int(rng.choice([50, 64, 100])), float(rng.choice([0.6, 0.8, 0.9, 0.99])), bool(rng.integers(0, 2))
        fo, fg = O.FrameData(k, d, 640, 480, u_right=ur), capi.FrameData(k, d, 640, 480, u_right=ur)
        a = O.search_by_projection(variant, fo, q, qd, blk, occ, th, ratio, ori)
        b = capi.search_by_projection(variant, fg, q, qd, blk, occ, th, ratio, ori)
        ok = a[0] == b[0] and all(np.array_equal(x, y) for x, y in zip(a[1:], b[1:]))
        return ok, f"SearchByProjection v{variant} n={n} nq={nq} r={radius} th={th} ratio={ratio} ori={ori} gate={ur is not None}"
    if kind == "init":
        n1, n2 = pick(rng), pick(rng)
        window, ratio, ori = int(rng.choice([10, 30, 100, 300])), float(rng.choice([0.7, 0.9, 0.99])), bool(rng.integers(0, 2))
        nbase = int(rng.choice([4, 64, 100000]))                # few distinct descriptors: lists fill with ties
        base = np.random.default_rng(t).integers(0, 256, (min(nbase, 512), 32), dtype=np.uint8)
        def mk(n, s):
            r = np.random.default_rng([seed, t, s])
            k = np.zeros(n, capi.KP_DTYPE)
            k["x"] = r.uniform(5, 635, n).astype(np.float32); k["y"] = r.uniform(5, 475, n).astype(np.float32)
            k["octave"] = r.integers(0, 3, n); k["angle"] = r.uniform(0, 360, n).astype(np.float32)
            dd = base[r.integers(0, len(base), n)].copy() if nbase < 100000 else r.integers(0, 256, (n, 32), dtype=np.uint8)
            if n:
                dd[np.arange(n), r.integers(0, 32, n)] ^= (1 << r.integers(0, 8, n)).astype(np.uint8)
            return k, dd
        (k1, d1), (k2, d2) = mk(n1, 1), mk(n2, 2)
        prev = (np.stack([k1["x"], k1["y"]], axis=1) + rng.normal(0, 3, (n1, 2))).astype(np.float32)
        a = O.search_for_initialization(O.FrameData(k1, d1, 640, 480), O.FrameData(k2, d2, 640, 480), prev, window, ratio, ori)
        b = capi.search_for_initialization(capi.FrameData(k1, d1, 640, 480), capi.FrameData(k2, d2, 640, 480), prev, window, ratio, ori)
        ok = a[0] == b[0] and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])
        return ok, f"SearchForInitialization n1={n1} n2={n2} window={window} ratio={ratio} ori={ori} base={nbase}"
    if kind == "bow":
        n = pick(rng, 1, 5000)
        A, B = make_pair(voc, n=n, seed=seed + t, levelsup=int(rng.integers(1, 4)))
        nb = pick(rng, 0, n)                                    # frames of different sizes
        B = tuple(x[:nb] for x in B) if rng.integers(0, 2) else B
        variant, ratio, ori = int(rng.integers(0, 2)), float(rng.choice([0.6, 0.75, 0.9])), bool(rng.integers(0, 2))
        (n_g, ma, mb), = capi.search_by_bow(variant, [(A, B)], nnratio=ratio, check_ori=ori)
        n_o, ma_o, mb_o = O.search_by_bow(variant, *A, *B, nnratio=ratio, check_ori=ori)
        return n_g == n_o and np.array_equal(ma, ma_o) and np.array_equal(mb, mb_o), f"SearchByBoW v{variant} nA={n} nB={len(B[0])} ratio={ratio} ori={ori}"
    if kind == "fuse":
        n, nq = pick(rng, 3), max(6, pick(rng, 0, 5000))
        variant = int(rng.integers(0, 3))
        kf, q, qd = make_fuse_case(seed * 31 + t, n=n, nq=nq, th=float(rng.choice([3.0, 4.0, 10.0])))
        bo, do = O.fuse(variant, kf, q, qd)
        bg, dg = capi.fuse(variant, _to_capi(kf), q, qd)
        return np.array_equal(bg, bo) and np.array_equal(dg, do), f"Fuse v{variant} n={n} nq={nq}"
    if kind == "tri":
        n1, n2 = pick(rng, 1), pick(rng, 1)
        a, b, F, epi = make_tri_pair(seed * 13 + t, n1=n1, n2=n2, mono_frac=float(rng.choice([0.0, 0.5, 1.0])), nodes=int(rng.choice([3, 24, 200])))
        only, ori = bool(rng.integers(0, 2)), bool(rng.integers(0, 2))
        nm, ms = capi.search_for_triangulation([(_to_capi(a), _to_capi(b))], [F], [epi], only, ori)
        n_o, m_o = O.search_for_triangulation(a, b, F, epi, only, ori)
        return nm[0] == n_o and np.array_equal(ms[0], m_o), f"SearchForTriangulation n1={n1} n2={n2} only_stereo={only} ori={ori}"
    sizes = rng.choice([0, 1, 2, 3, 7, 33, 64, 130, 500], int(rng.integers(1, 40)))
    offs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    dd = np.random.default_rng(t).integers(0, 256, (max(int(offs[-1]), 1), 32), dtype=np.uint8)
    dd[::2] = dd[0]
    ig, mg = capi.distinctive_descriptors(dd, offs)
    io, mo = O.distinctive_descriptors(dd, offs)
    return np.array_equal(ig, io) and np.array_equal(mg, mo), f"ComputeDistinctiveDescriptors groups={len(sizes)}"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--trials", type=int, default=700); ap.add_argument("--seed", type=int, default=1); ap.add_argument("--budget-s", type=float, default=300)
    args = ap.parse_args()
    voc = O.make_vocabulary(k=8, depth=4, seed=5)
    t0 = time.time()
    done = 0
    for t in range(args.trials):
        try:
            ok, desc = trial(t, args.seed, voc)
        except Exception as e:                                  # an error return of the library is a failure too
            ok, desc = False, f"trial raised {type(e).__name__}: {e}"
        if not ok:
            print(f"MISMATCH trial {t}: {desc}")
            sys.exit(1)
        done += 1
        if t % 25 == 0:
            print(f"ok trial {t}: {desc}", flush=True)
        if time.time() - t0 > args.budget_s:
            break
    print(f"MATCHER SOAK PASS: {done} trials, seed {args.seed}, {time.time() - t0:.0f} s")


if __name__ == "__main__":
    main()
