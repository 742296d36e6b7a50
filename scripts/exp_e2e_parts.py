"""attribute the e2e step time: which of H2D / kernels / D2H is exposed (run on a GPU box)"""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
from orb_slam2_map_b200 import capi, synth

F, W, H = 4096, 640, 480
CH = int(sys.argv[1]) if len(sys.argv) > 1 else 512
dev = torch.device("cuda:0")
seq = synth.frame_sequence(64, W, H, seed=1)
frames = torch.from_numpy(np.ascontiguousarray(np.tile(seq, (F // 64, 1, 1))))
host = torch.empty((F, H, W), dtype=torch.uint8, pin_memory=True); host.copy_(frames)
dimg = frames.to(dev)
ex = capi.Extractor(max_width=W, max_height=H, chunk_frames=CH, result_slots=F)
cap = ex.cap
pin = lambda shape, dt: torch.empty(shape, dtype=dt, pin_memory=True).numpy()
kp = pin((F, cap, 28), torch.uint8).view(capi.KP_DTYPE).reshape(F, cap); de = pin((F, cap, 32), torch.uint8)
cnt = pin((F,), torch.int32); m12 = pin((F - 1, cap), torch.int32); nm = pin((F - 1,), torch.int32)
hnp = host.numpy()

def timeit(name, fn, n=3):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / n * 1e3
    print(f"{name:50s} {ms:8.2f} ms  {F / ms * 1e3:9.0f} frames/s", flush=True)

import ctypes as C
def dev_only():
    ex.extract_batch(dimg.data_ptr(), nframes=F, width=W, height=H, frame_stride=W * H, row_stride=W, on_device=True)
    ex.search_init_pairs(0, F - 1, fetch=False); ex.sync()
timeit("device only (images resident, no outputs)", dev_only)
timeit("host in, counts only out, no match", lambda: ex.extract_batch_to_host(hnp, None, None, cnt))
timeit("host in, all extraction outputs, no match", lambda: ex.extract_batch_to_host(hnp, kp, de, cnt))
timeit("host in, match fused, counts + nmatches only", lambda: ex.extract_match_to_host(hnp, None, None, cnt, None, nm))
timeit("host in, match fused, all outputs", lambda: ex.extract_match_to_host(hnp, kp, de, cnt, m12, nm))
d = torch.empty((F, H, W), dtype=torch.uint8, device=dev)
timeit("H2D alone", lambda: d.copy_(host, non_blocking=True))
