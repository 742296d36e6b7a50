"""Randomised parity soak: liborbx.so (C ABI) vs the CPU oracle over random image sizes, extractor
parameters and image content.  Not collected by pytest; run on the GPU box:

    python scripts/soak_parity.py --trials 300 --seed 1 > gpurun_out/soak.log

Every trial checks, bit for bit: keypoints (x, y, octave, size, response, angle), descriptors, FAST
candidates per level, then SearchForInitialization between two consecutive frames, every fourth trial the
batch pipeline (orbx_extract_match_to_host, random chunk size and sequence length), and every eighth trial
ComputeStereoMatches on a synthetic stereo pair.  Exits non-zero at the first mismatch and prints the
trial's parameters so that it can be replayed with --only."""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

import oracle_lib as O                                   # noqa: E402
from orb_slam2_map_b200 import capi, synth               # noqa: E402


def box_blur(img, k):
    f = img.astype(np.float32)
    c = np.cumsum(np.cumsum(np.pad(f, ((k, 0), (k, 0))), axis=0), axis=1)
    s = c[k:, k:] - c[:-k, k:] - c[k:, :-k] + c[:-k, :-k]
    out = np.zeros_like(f)
    out[: s.shape[0], : s.shape[1]] = s / (k * k)
    return out


def content(kind, w, h, rng, seed):
    if kind == "canvas":
        return synth.textured_canvas(w, h, seed=seed)
    if kind == "noise":
        return rng.integers(0, 256, (h, w), dtype=np.uint8)
    if kind == "smooth":
        n = rng.integers(0, 256, (h, w)).astype(np.uint8)
        b = box_blur(n, int(rng.integers(2, 6)))
        b = (b - b.min()) / max(1e-6, b.max() - b.min()) * 255
        return b.astype(np.uint8)
    if kind == "lowcontrast":
        return (synth.textured_canvas(w, h, seed=seed) // 8 + 100).astype(np.uint8)
    if kind == "saturated":
        c = synth.textured_canvas(w, h, seed=seed).astype(np.int32)
        return np.clip((c - 128) * 4 + 128, 0, 255).astype(np.uint8)
    if kind == "sparse":
        img = np.full((h, w), int(rng.integers(0, 256)), np.uint8)
        for _ in range(int(rng.integers(1, 60))):
            y, x = int(rng.integers(0, h - 6)), int(rng.integers(0, w - 6))
            img[y:y + int(rng.integers(1, 7)), x:x + int(rng.integers(1, 7))] = int(rng.integers(0, 256))
        return img
    raise ValueError(kind)


KINDS = ["canvas", "canvas", "canvas", "noise", "smooth", "lowcontrast", "saturated", "sparse"]
FIELDS = ("x", "y", "size", "octave", "class_id", "response", "angle")


def trial(t, seed):
    rng = np.random.default_rng([seed, t])
    w = int(rng.integers(96, 1000))
    h = int(rng.integers(96, 760))
    # round(width / height) == 0 is out of the reference's domain (DistributeOctTree divides by it; oracle and
    # kernels both return no keypoints): keep a few such shapes, turn most of them sideways
    if h > 1.8 * w and rng.integers(0, 5) > 0:
        w, h = h, w
    nf = int(rng.choice([100, 300, 500, 1000, 1500, 2000, 3000]))
    scale = float(rng.choice([1.2, 1.2, 1.2, 1.1, 1.25, 1.5, 2.0]))
    nlevels = int(rng.choice([8, 8, 8, 4, 6, 1, 10, 12]))
    # the smallest level must stay larger than the 2*EDGE_THRESHOLD border (the reference assumes it too)
    while nlevels > 1 and min(w, h) / scale ** (nlevels - 1) < 60:
        nlevels -= 1
    ini_th, min_th = [(20, 7), (20, 7), (12, 5), (30, 10), (40, 3), (7, 7)][int(rng.integers(0, 6))]
    kind = KINDS[int(rng.integers(0, len(KINDS)))]
    desc = f"trial {t}: {w}x{h} nf={nf} scale={scale} nlevels={nlevels} th=({ini_th},{min_th}) content={kind}"
    a = content(kind, w, h, rng, seed * 100003 + t)
    shift = int(rng.integers(1, 9))
    b = np.roll(a, (shift // 2, shift), axis=(0, 1))
    ex = capi.Extractor(nfeatures=nf, scale=scale, nlevels=nlevels, ini_th=ini_th, min_th=min_th, max_width=w, max_height=h)
    orc = O.Extractor(nfeatures=nf, scale=scale, nlevels=nlevels, ini_th=ini_th, min_th=min_th)
    res = []
    for img in (a, b):
        ko, do = orc.extract(img)
        kg, dg = ex.extract(img)
        for l in range(nlevels):
            cg, co = ex.candidates(0, l), orc.candidates(l)
            if cg.shape != co.shape or not np.array_equal(cg, co):
                return False, desc + f" -> FAST candidates differ at level {l}", 0
        if len(kg) != len(ko):
            return False, desc + f" -> keypoint count {len(kg)} vs {len(ko)}", 0
        for f in FIELDS:
            if not np.array_equal(kg[f], ko[f]):
                return False, desc + f" -> keypoint field {f} differs", 0
        if not np.array_equal(dg, do):
            return False, desc + " -> descriptors differ", 0
        res.append((ko, do))
    (k1, d1), (k2, d2) = res
    nkp = len(k1)
    # SearchForInitialization wants level-0 features on both sides
    window = int(rng.choice([100, 30, 250, 10]))
    nnratio = float(rng.choice([0.9, 0.7, 0.95]))
    ori = bool(rng.integers(0, 2))
    prev = np.stack([k1["x"], k1["y"]], axis=1).astype(np.float32)
    if len(k1) and len(k2):
        n_o, m_o, p_o, _ = O.search_for_initialization(O.FrameData(k1, d1, w, h), O.FrameData(k2, d2, w, h), prev, window, nnratio, ori)
        n_g, m_g, p_g = capi.search_for_initialization(capi.FrameData(k1, d1, w, h), capi.FrameData(k2, d2, w, h), prev, window, nnratio, ori)
        if n_g != n_o or not np.array_equal(m_g, m_o) or not np.array_equal(p_g, p_o):
            return False, desc + f" -> SearchForInitialization(window={window}, nnratio={nnratio}, ori={ori}) differs", nkp
    if t % 4 == 1:
        # the batch pipeline (ramped / tapered chunks, matcher folded in) on a sequence built from the two frames
        n = int(rng.integers(2, 12))
        pick = rng.integers(0, 2, n)
        frames = np.stack([a if p == 0 else b for p in pick])
        exb = capi.Extractor(nfeatures=nf, scale=scale, nlevels=nlevels, ini_th=ini_th, min_th=min_th, max_width=w, max_height=h,
                             chunk_frames=int(rng.choice([1, 2, 3, 5, 16])), result_slots=n)
        cap = exb.cap
        kp = np.zeros((n, cap), capi.KP_DTYPE); de = np.zeros((n, cap, 32), np.uint8); cnt = np.zeros(n, np.int32)
        m12 = np.full((n - 1, cap), -7, np.int32); nm = np.full(n - 1, -7, np.int32)
        exb.extract_match_to_host(frames, kp, de, cnt, m12, nm, window, nnratio, ori)
        for f in range(n):
            ko, do = res[pick[f]]
            if cnt[f] != len(ko) or kp[f, :cnt[f]].tobytes() != np.ascontiguousarray(ko).tobytes() or not np.array_equal(de[f, :cnt[f]], do):
                return False, desc + f" -> batch pipeline frame {f} of {n} differs from the single-frame result", nkp
        for f in range(n - 1):
            (ka, da), (kb, db) = res[pick[f]], res[pick[f + 1]]
            if len(ka) == 0:
                continue
            pv = np.stack([ka["x"], ka["y"]], axis=1).astype(np.float32)
            n_o, m_o, _, _ = O.search_for_initialization(O.FrameData(ka, da, w, h), O.FrameData(kb, db, w, h), pv, window, nnratio, ori)
            if nm[f] != n_o or not np.array_equal(m12[f, :cnt[f]], m_o):
                return False, desc + f" -> batch pipeline pair {f} of {n - 1} differs", nkp
        exb.close()
    if t % 8 == 0 and nlevels == 8 and scale == 1.2 and w >= 200:
        left, right = synth.stereo_pair(w, h, seed=seed + t)
        el, er = O.Extractor(nf), O.Extractor(nf)
        kl, dl = el.extract(left)
        kr, dr = er.extract(right)
        pl = [el.level(l) for l in range(8)]
        pr = [er.level(l) for l in range(8)]
        tb = el.tables()
        bf, fx = 386.1448, 718.856
        u_o, z_o, c_o = O.compute_stereo_matches(kl, dl, kr, dr, pl, pr, tb["scale"], tb["inv_scale"], bf, fx)
        u_g, z_g, c_g = capi.compute_stereo_matches(kl, dl, kr, dr, pl, pr, tb["scale"], tb["inv_scale"], bf, fx)
        if c_g != c_o or not np.array_equal(u_g, u_o) or not np.array_equal(z_g, z_o):
            return False, desc + " -> ComputeStereoMatches differs", nkp
    ex.close()
    return True, desc, nkp


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--trials", type=int, default=200)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--only", type=int, default=-1)
    ap.add_argument("--budget-s", type=float, default=600.0)
    args = ap.parse_args()
    t0 = time.time()
    done = 0
    kps = 0
    ts = [args.only] if args.only >= 0 else range(args.trials)
    for t in ts:
        ok, desc, n = trial(t, args.seed)
        if not ok:
            print("MISMATCH", desc, flush=True)
            sys.exit(1)
        done += 1
        kps += n
        if t % 10 == 0 or args.only >= 0:
            print("ok", desc, f"kp={n}", flush=True)
        if time.time() - t0 > args.budget_s:
            print(f"time budget reached after {done} trials")
            break
    print(f"SOAK PASS: {done} trials, seed {args.seed}, {kps} keypoints compared, {time.time() - t0:.0f} s")


if __name__ == "__main__":
    main()
