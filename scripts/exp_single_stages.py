"""Per-stage device time of ONE 640x480 frame (plain launches, CUDA events placed by liborbx around every stage)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from orb_slam2_map_b200 import capi, synth
W, H = 640, 480
dev = torch.device("cuda", 0)
frames = synth.frame_sequence_torch(8, W, H, seed=0, device=dev)
stream = torch.cuda.Stream(device=dev); torch.cuda.set_stream(stream)
ex = capi.Extractor(max_width=W, max_height=H, chunk_frames=1, result_slots=1, device=0, stream=stream.cuda_stream)
run = lambda i: ex.extract_batch(frames[i % 8].data_ptr(), nframes=1, width=W, height=H, frame_stride=W * H, row_stride=W, on_device=True)
for i in range(50): run(i)
ex.sync(); ex.profile(); ex.set_profiling(True)
N = 200
for i in range(N): run(i)
ms, launches = ex.profile()
print({k: round(v / N * 1e3, 2) for k, v in ms.items() if launches[k]}, "us per frame; sum", round(sum(ms.values()) / N * 1e3, 1))
