"""Latency of ONE RGB-D frame through the frame front (orbx_frames_rgbd, nframes = 1) + downloads: where the
single-frame graph path of orbx_extract is not used yet."""
import os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from orb_slam2_map_b200 import capi, synth
W, H = 640, 480
gray = synth.frame_sequence(8, W, H, seed=0)
rng = np.random.default_rng(0)
color = np.stack([np.clip(gray.astype(np.int16) + rng.integers(-12, 13, gray.shape), 0, 255).astype(np.uint8) for _ in range(3)], axis=-1)
depth = rng.integers(500, 30000, gray.shape).astype(np.uint16)
cam = capi.Camera(517.306408, 516.469215, 318.643040, 255.313989, 0.262383, -0.953104, -0.005358, 0.002628, 1.163314, 40.0)
ex = capi.Extractor(max_width=W, max_height=H, chunk_frames=1, result_slots=1, device=0)
def one(i):
    ex.frames_rgbd(color[i % 8][None], depth[i % 8][None], cam, True, 1.0 / 5000.0, 0)
    ex.download(0, 1); ex.download_frame_front(0, 1)
for i in range(20): one(i)
ts = []
for i in range(200):
    t0 = time.perf_counter(); one(i); ts.append(time.perf_counter() - t0)
print(json.dumps({"rgbd_single_frame_us_median": float(np.median(ts)) * 1e6, "note": "orbx_frames_rgbd(nframes=1) + orbx_download + orbx_download_frame_front through the ctypes wrapper"}))
