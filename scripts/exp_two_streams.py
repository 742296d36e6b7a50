"""Experiment: does running two independent chunk pipelines concurrently (two handles, two streams, two host
threads) raise extraction throughput?  Compares against one handle on all frames."""
import sys, time, threading
sys.path.insert(0, ".")
import torch
from orb_slam2_map_b200 import capi, synth

W, H, F = 640, 480, 4096
dev = torch.device("cuda", 0)
frames = synth.frame_sequence_torch(F, W, H, seed=0, device=dev)
PARAMS = dict(nfeatures=1000, scale=1.2, nlevels=8, ini_th=20, min_th=7)


def run(nlanes, chunk, steps=3):
    per = F // nlanes
    streams = [torch.cuda.Stream(device=dev) for _ in range(nlanes)]
    exs = [capi.Extractor(**PARAMS, max_width=W, max_height=H, chunk_frames=chunk, result_slots=per, device=0,
                          stream=s.cuda_stream) for s in streams]

    def work(i, n):
        for _ in range(n):
            exs[i].extract_batch(frames.data_ptr() + i * per * W * H, nframes=per, width=W, height=H, frame_stride=W * H,
                                 row_stride=W, on_device=True, first_slot=0)
            exs[i].search_init_pairs(0, per - 1, 100, 0.9, True, fetch=False)

    def all_lanes(n):
        th = [threading.Thread(target=work, args=(i, n)) for i in range(nlanes)]
        [t.start() for t in th]; [t.join() for t in th]
        torch.cuda.synchronize()

    all_lanes(2)
    t0 = time.perf_counter()
    all_lanes(steps)
    dt = time.perf_counter() - t0
    for e in exs:
        e.close()
    return F * steps / dt


for nl, ch in [(1, 512), (2, 256), (2, 512), (3, 256), (4, 128), (4, 256)]:
    print(f"lanes {nl} chunk {ch}: {run(nl, ch):,.0f} frames/s", flush=True)
