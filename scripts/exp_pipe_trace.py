"""print the chunk timeline of the host-image pipeline (ORBX_PIPE_TRACE) on a GPU box"""
import os, sys
import numpy as np, torch
sys.path.insert(0, ".")
from orb_slam2_map_b200 import capi, synth
F, W, H = 4096, 640, 480
CH = int(sys.argv[1]) if len(sys.argv) > 1 else 512
seq = synth.frame_sequence(64, W, H, seed=1)
host = torch.empty((F, H, W), dtype=torch.uint8, pin_memory=True); host.copy_(torch.from_numpy(np.ascontiguousarray(np.tile(seq, (F // 64, 1, 1)))))
ex = capi.Extractor(max_width=W, max_height=H, chunk_frames=CH, result_slots=F)
cap = ex.cap
pin = lambda shape, dt: torch.empty(shape, dtype=dt, pin_memory=True).numpy()
kp = pin((F, cap, 28), torch.uint8).view(capi.KP_DTYPE).reshape(F, cap); de = pin((F, cap, 32), torch.uint8)
cnt = pin((F,), torch.int32); m12 = pin((F - 1, cap), torch.int32); nm = pin((F - 1,), torch.int32)
for _ in range(2): ex.extract_match_to_host(host.numpy(), kp, de, cnt, m12, nm)
os.environ["ORBX_PIPE_TRACE"] = "1"
ex.extract_match_to_host(host.numpy(), kp, de, cnt, m12, nm)
