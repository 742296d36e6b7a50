"""Host->device copy bandwidth of every GPU alone, of every pair copying at once and of all GPUs at once (pinned host memory,
one process, one stream per GPU, CUDA events): which GPUs share a host bridge, and what the box's aggregate H2D rate is.
Prints one JSON object."""
import itertools, json, sys
import torch

n = torch.cuda.device_count()
MB = 256
host = [torch.empty(MB << 20, dtype=torch.uint8, pin_memory=True) for _ in range(n)]
dev = [torch.empty(MB << 20, dtype=torch.uint8, device=f"cuda:{i}") for i in range(n)]
streams = [torch.cuda.Stream(device=i) for i in range(n)]


def run(gpus, reps=4):
    ev = {}
    for g in gpus:
        torch.cuda.synchronize(g)
    for g in gpus:
        with torch.cuda.device(g), torch.cuda.stream(streams[g]):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                dev[g].copy_(host[g], non_blocking=True)
            e1.record()
            ev[g] = (e0, e1)
    out = {}
    for g in gpus:
        torch.cuda.synchronize(g)
        out[g] = reps * MB / 1024 / (ev[g][0].elapsed_time(ev[g][1]) * 1e-3)
    return out


for g in range(n):
    run([g], 1)
alone = {g: run([g])[g] for g in range(n)}
pairs = {}
for a, b in itertools.combinations(range(n), 2):
    r = run([a, b])
    pairs[f"{a},{b}"] = [round(r[a], 1), round(r[b], 1)]
allg = run(list(range(n)))
halves = {}
if n >= 4:
    lo, hi = list(range(n // 2)), list(range(n // 2, n))
    halves = {"first_half": round(sum(run(lo).values()), 1), "second_half": round(sum(run(hi).values()), 1)}
print(json.dumps({"gpus": n, "alone_gib_s": {str(k): round(v, 1) for k, v in alone.items()}, "pairs_gib_s": pairs,
                  "all_at_once_gib_s": {str(k): round(v, 1) for k, v in allg.items()}, "all_at_once_sum": round(sum(allg.values()), 1), **halves}))
