#!/usr/bin/env python
"""Regenerates the golden fixtures in this directory.  Run from the repo root IN THE BUILD
CONTAINER (needs the cv2 4.13.0 wheel and /root/reference for oracle/_ref):

    python tests/golden/make_golden.py

* cv2_primitives.npz  — outputs of the REAL OpenCV 4.13.0 (cv2 wheel) for the five primitives the
  reference calls (resize INTER_LINEAR 8U, GaussianBlur 7x7 s=2 REFLECT_101, copyMakeBorder
  REFLECT_101, FAST 9/16 + NMS at th=7/20, fastAtan2) on seeded inputs.  Pins the oracle's
  integer/float models.
* ref_orbextractor.npz — outputs of the UNMODIFIED reference src/ORBextractor.cc (oracle/_ref:
  compiled over oracle/ref_shim with those primitives and a bump allocator) on seeded frames:
  full keypoint+descriptor bytes for small cases, SHA-256 for the BASELINE.json shapes.

Inputs are NOT stored: they are regenerated from seeds with orb_slam2_map_b200.synth
(splitmix64 counter hashes — identical bytes on any machine).
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from orb_slam2_map_b200 import synth  # noqa: E402

RESIZE_CASES = [(97, 61, 81, 51), (160, 120, 133, 100), (64, 48, 53, 40), (214, 161, 179, 134), (31, 17, 26, 14)]
RESIZE_HASH_CASES = [(640, 480, 533, 400), (533, 400, 444, 333), (1241, 376, 1034, 313), (1920, 1080, 1600, 900)]
BLUR_CASES = [(97, 61), (64, 48), (179, 134), (9, 8), (7, 40)]
BLUR_HASH_CASES = [(640, 480), (1241, 376)]
FAST_CASES = [(37, 37, 1), (37, 37, 2), (36, 38, 3), (43, 40, 4), (160, 120, 5), (12, 9, 6), (7, 7, 7)]
ATAN_N = 4096


def noise_image(w, h, seed):
    """iid uniform bytes (worst case for the fixed-point models)."""
    n = w * h
    ctr = np.arange((n + 7) // 8, dtype=np.uint64) + (np.uint64(seed) << np.uint64(36))
    return synth.splitmix64(ctr).view(np.uint8)[:n].reshape(h, w).copy()


def test_image(w, h, seed):
    """half textured canvas, half noise: corners at every threshold"""
    if w >= 24 and h >= 24:
        img = synth.textured_canvas(w, h, seed=seed)
        nz = noise_image(w, h, seed + 100)
        img[:, : w // 3] = nz[:, : w // 3]
        return img
    return noise_image(w, h, seed)


def atan_inputs():
    h = synth.splitmix64(np.arange(2 * ATAN_N, dtype=np.uint64) + np.uint64(77 << 40))
    v = (h % np.uint64(400001)).astype(np.int64) - 200000
    y, x = v[:ATAN_N].astype(np.float32), v[ATAN_N:].astype(np.float32)
    y[:8] = [0, 0, 1, -1, 5, -5, 0, 3]
    x[:8] = [0, 1, 0, 0, 5, -5, -7, -3]
    return y, x


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


EXTRACT_FULL = [  # (w, h, nfeatures, scale, nlevels, ini, min, seed)
    (320, 240, 300, 1.2, 8, 20, 7, 11),
    (200, 150, 150, 1.3, 5, 25, 9, 12),
]
EXTRACT_HASH = [
    (640, 480, 1000, 1.2, 8, 20, 7, 0),
    (640, 480, 1000, 1.2, 8, 20, 7, 1),
    (1241, 376, 2000, 1.2, 8, 20, 7, 2),
    (1920, 1080, 5000, 1.2, 8, 20, 7, 3),
    (752, 480, 1200, 1.2, 8, 20, 7, 4),
]


def extract_input(w, h, seed):
    return synth.textured_canvas(w, h, seed=seed)


def make_cv2():
    import cv2
    assert cv2.__version__.startswith("4.13"), cv2.__version__
    out = {"cv2_version": np.array(cv2.__version__)}
    for i, (sw, sh, dw, dh) in enumerate(RESIZE_CASES):
        out[f"resize_{i}"] = cv2.resize(test_image(sw, sh, 10 + i), (dw, dh), interpolation=cv2.INTER_LINEAR)
    out["resize_hash"] = np.array([sha(cv2.resize(test_image(sw, sh, 20 + i), (dw, dh), interpolation=cv2.INTER_LINEAR))
                                   for i, (sw, sh, dw, dh) in enumerate(RESIZE_HASH_CASES)])
    for i, (w, h) in enumerate(BLUR_CASES):
        out[f"blur_{i}"] = cv2.GaussianBlur(test_image(w, h, 30 + i), (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
        out[f"border_{i}"] = cv2.copyMakeBorder(test_image(w, h, 30 + i), 5, 5, 5, 5, cv2.BORDER_REFLECT_101) if min(w, h) > 5 \
            else cv2.copyMakeBorder(test_image(w, h, 30 + i), 3, 3, 3, 3, cv2.BORDER_REFLECT_101)
    out["blur_hash"] = np.array([sha(cv2.GaussianBlur(test_image(w, h, 40 + i), (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101))
                                 for i, (w, h) in enumerate(BLUR_HASH_CASES)])
    for i, (w, h, seed) in enumerate(FAST_CASES):
        img = test_image(w, h, 50 + seed)
        for th in (7, 20):
            det = cv2.FastFeatureDetector_create(th, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
            kps = det.detect(img)
            out[f"fast_{i}_{th}"] = np.array([[k.pt[0], k.pt[1], k.response] for k in kps], np.float32).reshape(-1, 3)
        det = cv2.FastFeatureDetector_create(20, False, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
        out[f"fast_nonms_{i}"] = np.array([[k.pt[0], k.pt[1]] for k in det.detect(img)], np.float32).reshape(-1, 2)
    # a strided ROI view must give the same result as its copy (the reference passes ROIs)
    big = test_image(160, 120, 55)
    roi = big[16:53, 40:77]
    det = cv2.FastFeatureDetector_create(7, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    out["fast_roi"] = np.array([[k.pt[0], k.pt[1], k.response] for k in det.detect(roi)], np.float32).reshape(-1, 3)
    y, x = atan_inputs()
    out["atan2"] = np.array([cv2.fastAtan2(float(a), float(b)) for a, b in zip(y, x)], np.float32)
    np.savez_compressed(os.path.join(HERE, "cv2_primitives.npz"), **out)
    print("wrote cv2_primitives.npz", os.path.getsize(os.path.join(HERE, "cv2_primitives.npz")), "bytes")


def make_ref():
    import ref_lib as R
    assert R.available(), "build oracle/_ref first: make -C oracle ref"
    out = {}
    for i, (w, h, nf, sc, nl, it, mt, seed) in enumerate(EXTRACT_FULL):
        r = R.RefExtractor(nf, sc, nl, it, mt)
        k, d = r.extract(extract_input(w, h, seed))
        out[f"full_kps_{i}"] = k.view(np.uint8).reshape(len(k), 28)
        out[f"full_desc_{i}"] = d
        out[f"full_level3_border_{i}"] = r.level(min(3, nl - 1), border=True)
    hashes = []
    for (w, h, nf, sc, nl, it, mt, seed) in EXTRACT_HASH:
        r = R.RefExtractor(nf, sc, nl, it, mt)
        k, d = r.extract(extract_input(w, h, seed))
        hashes.append([str(len(k)), sha(k), sha(d), sha(np.concatenate([r.level(l).ravel() for l in range(nl)]))])
    out["hash"] = np.array(hashes)
    t = R.RefExtractor().tables()
    for k, v in t.items():
        out["table_" + k] = v
    np.savez_compressed(os.path.join(HERE, "ref_orbextractor.npz"), **out)
    print("wrote ref_orbextractor.npz", os.path.getsize(os.path.join(HERE, "ref_orbextractor.npz")), "bytes")


if __name__ == "__main__":
    make_cv2()
    make_ref()
