#!/usr/bin/env python
"""Golden vectors for the RGB-D frame front, generated with the cv2 4.13.0 wheel:
cvtColor (RGB/BGR/RGBA/BGRA -> GRAY, 8U) and undistortPoints (K, D = TUM1 and a strong fisheye-like case,
R = I, P = K) on seeded inputs.  Run once where cv2 is importable; commits tests/golden/frame_front.npz."""
import os

import cv2
import numpy as np

rng = np.random.default_rng(20261104)
out = {"cv2_version": np.array(cv2.__version__)}
rgb = rng.integers(0, 256, (48, 67, 3), dtype=np.uint8)
rgba = rng.integers(0, 256, (31, 50, 4), dtype=np.uint8)
out["rgb"] = rgb
out["rgba"] = rgba
out["gray_rgb"] = cv2.cvtColor(rgb, cv2.COLOR_RGB2GRAY)
out["gray_bgr"] = cv2.cvtColor(rgb, cv2.COLOR_BGR2GRAY)
out["gray_rgba"] = cv2.cvtColor(rgba, cv2.COLOR_RGBA2GRAY)
out["gray_bgra"] = cv2.cvtColor(rgba, cv2.COLOR_BGRA2GRAY)
# Examples/RGB-D/TUM1.yaml:9-19 and a stronger distortion
cams = np.array([[517.306408, 516.469215, 318.643040, 255.313989], [420.0, 415.0, 330.5, 236.25]], np.float32)
dists = np.array([[0.262383, -0.953104, -0.005358, 0.002628, 1.163314], [-0.31, 0.12, 0.0011, -0.0007, -0.02]], np.float32)
pts = (rng.random((1500, 2)) * [640, 480]).astype(np.float32)
pts[:4] = [[0, 0], [640, 0], [0, 480], [640, 480]]
out["cams"] = cams
out["dists"] = dists
out["pts"] = pts
for i in range(2):
    K = np.array([[cams[i, 0], 0, cams[i, 2]], [0, cams[i, 1], cams[i, 3]], [0, 0, 1]], np.float32)
    out[f"undist{i}"] = cv2.undistortPoints(pts.reshape(-1, 1, 2), K, dists[i], None, K).reshape(-1, 2)
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "frame_front.npz"), **out)
print({k: getattr(v, "shape", None) for k, v in out.items()})
