#!/usr/bin/env python
"""bench.py — headline benchmark of the B200-native ORB front end.

Workload (BASELINE.json configs[1]): a batch of 4096 synthetic 640x480 frames; one "step" =
ORBextractor::operator() on every frame (nFeatures 1000, 8 levels, 1.2, FAST 20/7) + consecutive-
frame ORBmatcher::SearchForInitialization (window 100, nnratio 0.9, orientation check) on the
4095 pairs.  Metric: frames/s.

  value     device-resident: frames already in HBM, results left in HBM, CUDA events on the
            stream the kernels run on.
  e2e       same step through the C ABI with HOST (pinned) images in and HOST results out:
            H2D of the frames and D2H of keypoints, descriptors, counts and matches are inside the
            timed region.
  roofline  dominant kernel (largest share of device time, measured live with CUDA events placed
            around every stage by liborbx), algorithmic bytes / its time vs MEASURED_PEAKS.json.
  cpu_baseline  the CPU oracle (restatement of the reference's CPU path) on the box's host cores,
            bounded sample of the same frames.

`--workload single` is BASELINE configs[0]: the latency of ONE frame per orbx_extract call (host image in, host
keypoints + descriptors out) next to the reference's ORBextractor.cc on one host thread; `--workload hd|stereo|knn|rgbd`
are configs[3], [2], [4] and the RGB-D frame front.

`--impl reference` times the reference's CPU path (oracle/_ref when built from the reference's own
ORBextractor.cc, else the oracle port) with all host threads on the same workload.
Multi-GPU: one process per GPU (torchrun), frames sharded by rank, no data-path collective
(weak scaling: every rank processes its own 4096-frame batch).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H = 640, 480
PARAMS = dict(nfeatures=1000, scale=1.2, nlevels=8, ini_th=20, min_th=7)
WINDOW, NNRATIO = 100, 0.9
METRIC = "ORB frames/s (640x480, 1000 kp, extract+match)"

# algorithmic bytes per 640x480 frame (DESIGN.md "Roofline")
def level_sizes(w, h, nlevels=8, scale=1.2):
    sf = np.float32(1.0)
    out = []
    for l in range(nlevels):
        inv = np.float32(1.0) / sf
        out.append((int(np.rint(np.float32(w) * inv)), int(np.rint(np.float32(h) * inv))))
        sf = np.float32(np.float64(sf) * np.float64(np.float32(scale)))
    return out


def algorithmic_bytes(w, h, nkp):
    px = [a * b for a, b in level_sizes(w, h)]
    b_resize = sum(px[:-1]) + sum(px[1:])             # level l reads raw_{l-1}, writes raw_l
    b_blur = 2 * sum(px)                              # reads raw_l, writes blur_l
    b_pyr = px[0] + sum(px[:-1]) + px[0] + 2 * sum(px[1:])   # fused formulation (the graded figure)
    b_fast = sum(px)
    b_desc = nkp * (749 + 512 + 32 + 28)
    return dict(resize=b_resize, blur=b_blur, fast=b_fast, describe=b_desc, pyr_fused=b_pyr,
                total=b_pyr + b_fast + b_desc)


def ncu_constants():
    """Per-kernel dram bytes and warp instructions per 640x480 frame from ONE `ncu --set full` capture of the 512-frame
    chunk the bench launches (profiles/ncu_constants.json, written by scripts/ncu_constants.py from the .ncu-rep).  The file
    names the kernel sources it was captured on (sha256 of orb_slam2_map_b200/csrc/*.cu,*.h,*.cuh): when they changed
    since, the numbers are stale and `traffic` / `issue` are reported as null rather than quoted."""
    path = os.path.join(ROOT, "profiles", "ncu_constants.json")
    try:
        with open(path) as f:
            c = json.load(f)
    except Exception:
        return None, "profiles/ncu_constants.json missing"
    if c.get("csrc_sha256") != csrc_digest():
        return None, f"profiles/ncu_constants.json was captured on other kernel sources ({c.get('csrc_sha256', '?')[:12]})"
    return c, None


def csrc_digest():
    import hashlib
    d = os.path.join(ROOT, "orb_slam2_map_b200", "csrc")
    h = hashlib.sha256()
    for name in sorted(os.listdir(d)):
        if name.endswith((".cu", ".cuh", ".h")):
            with open(os.path.join(d, name), "rb") as f:
                h.update(name.encode()); h.update(f.read())
    return h.hexdigest()


def bind_to_gpu_numa(local):
    """Pin this rank to the CPUs next to its GPU (sysfs local_cpulist of the GPU's PCI function) BEFORE any pinned
    host buffer is allocated, so the end-to-end copies of N ranks do not all cross one socket.  Returns a small
    dict for the JSON line; silently does nothing where the box exposes no topology."""
    info = {"bound": False}
    try:
        out = subprocess.run(["nvidia-smi", "--query-gpu=pci.bus_id", "--format=csv,noheader", "-i", str(local)],
                             capture_output=True, text=True, timeout=20).stdout.strip().splitlines()[0].strip().lower()
        bdf = out[-12:] if len(out) >= 12 else out                  # 00000000:1b:00.0 -> 0000:1b:00.0
        base = "/sys/bus/pci/devices/" + bdf
        node = int(open(base + "/numa_node").read().strip())
        cpus = set()
        for part in open(base + "/local_cpulist").read().strip().split(","):
            if not part:
                continue
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        avail = os.sched_getaffinity(0)
        cpus &= avail
        info.update({"gpu_pci": bdf, "numa_node": node, "local_cpus": len(cpus)})
        if cpus and len(cpus) < len(avail):
            os.sched_setaffinity(0, cpus)
            info["bound"] = True
    except Exception as e:      # no topology in this container / VM
        info["note"] = type(e).__name__
    return info


_SAVED_STDOUT = None


def quiet_stdout():
    """Libraries (NCCL's version banner under torchrun) write to file descriptor 1; the contract is ONE JSON line on
    stdout.  Everything written to fd 1 before the line goes to stderr instead; emit() restores the descriptor."""
    global _SAVED_STDOUT
    if _SAVED_STDOUT is None:
        sys.stdout.flush()
        _SAVED_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    """The ONE line goes to the ORIGINAL stdout; descriptor 1 itself stays pointed at stderr for the rest of the process, so
    that whatever a library still prints afterwards (NCCL's communicator teardown under NCCL_DEBUG=INFO) cannot follow it."""
    sys.stdout.flush()
    data = (json.dumps(line) + "\n").encode()
    if _SAVED_STDOUT is not None:
        os.write(_SAVED_STDOUT, data)
    else:
        os.write(1, data)


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks/throttle reasons during the timed region."""

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "100"], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def oracle_mod():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib  # test infrastructure: only the cpu_baseline / reference legs use it
    return oracle_lib


def cpu_level():
    """highest x86-64 micro-architecture level of this host among the ones oracle/Makefile builds (v4 = AVX-512, v3 = AVX2+FMA)"""
    try:
        flags = set(open("/proc/cpuinfo").read().split("flags", 1)[1].split("\n", 1)[0].split())
    except Exception:
        return None
    if {"avx512f", "avx512bw", "avx512cd", "avx512dq", "avx512vl"} <= flags:
        return "v4"
    if {"avx2", "fma", "bmi2", "movbe"} <= flags:
        return "v3"
    return None


def ref_lib(fast=False):
    """oracle/_ref: the UNMODIFIED reference ORBextractor.cc over the cv shim, if it was built.  fast: the -O3 -march build
    ("as the reference compiles it", timing only) for this host's level, else None."""
    name = "libref_orbextractor.so"
    if fast:
        lvl = cpu_level()
        if lvl is None:
            return None
        name = f"libref_orbextractor_o3{lvl}.so"
    p = os.path.join(ROOT, "oracle", "_ref", name)
    if not os.path.exists(p):
        return None
    L = C.CDLL(p)
    L.ref_extract_match_many.restype = C.c_int64
    L.ref_extractor_create.restype = C.c_void_p
    return L


def cpu_single_thread_ms(frames: np.ndarray, fast=False):
    """Per-frame times (ms) of ORBextractor::operator() of oracle/_ref on ONE thread, one call per frame — the way ORB-SLAM2
    runs it (Frame::ExtractORB, src/Frame.cc:247-253)."""
    L = ref_lib(fast)
    if L is None:
        return None
    O = oracle_mod()
    p = O.Params(PARAMS["nfeatures"], PARAMS["scale"], PARAMS["nlevels"], PARAMS["ini_th"], PARAMS["min_th"])
    e = C.c_void_p(L.ref_extractor_create(C.byref(p)))
    cap = PARAMS["nfeatures"] + 64 * PARAMS["nlevels"] + 64
    kps = np.zeros(cap, O.KP_DTYPE); desc = np.zeros((cap, 32), np.uint8)
    out = []
    for i in range(len(frames)):
        f = frames[i]
        t0 = time.perf_counter()
        L.ref_extract(e, f.ctypes.data_as(C.c_void_p), C.c_int(f.shape[1]), C.c_int(f.shape[0]), C.c_size_t(f.strides[0]),
                      kps.ctypes.data_as(C.c_void_p), desc.ctypes.data_as(C.c_void_p), C.c_int(cap))
        out.append((time.perf_counter() - t0) * 1e3)
    L.ref_extractor_destroy(e)
    return out


def cv2_primitives_ms(frames: np.ndarray):
    """Single-thread time per frame of the REAL OpenCV primitives the reference calls (SIMD build of the cv2 wheel, cv2.setNumThreads(1)):
    7 x cv::resize, 8 x GaussianBlur(7x7, 2) and one cv::FAST call per 30-px grid cell at iniThFAST (+ minThFAST on empty cells), on
    the same frames: the OpenCV-dependent part of ORBextractor::operator() when the reference is linked against OpenCV instead of
    the scalar models behind oracle/ref_shim.  Excludes the quadtree, IC_Angle and the descriptors (the reference's own C++)."""
    try:
        import cv2
    except Exception:
        return None
    cv2.setNumThreads(1)
    sizes = level_sizes(W, H, PARAMS["nlevels"], PARAMS["scale"])
    det_ini = cv2.FastFeatureDetector_create(PARAMS["ini_th"], True)
    det_min = cv2.FastFeatureDetector_create(PARAMS["min_th"], True)
    out = []
    for f in frames:
        t0 = time.perf_counter()
        levels = [f]
        for (w, h) in sizes[1:]:
            levels.append(cv2.resize(levels[-1], (w, h), interpolation=cv2.INTER_LINEAR))
        t1 = time.perf_counter()
        for lv in levels:
            cv2.GaussianBlur(lv, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
        t2 = time.perf_counter()
        ncell = 0
        for lv in levels:
            h, w = lv.shape
            minb, maxbx, maxby = 16, w - 16, h - 16
            ncols, nrows = int((maxbx - minb) / 30), int((maxby - minb) / 30)
            if ncols <= 0 or nrows <= 0:
                continue
            wc, hc = -(-(maxbx - minb) // ncols), -(-(maxby - minb) // nrows)
            for i in range(nrows):
                y0 = minb + i * hc
                if y0 >= maxby - 3:
                    continue
                y1 = min(y0 + hc + 6, maxby)
                for j in range(ncols):
                    x0 = minb + j * wc
                    if x0 >= maxbx - 6:
                        continue
                    x1 = min(x0 + wc + 6, maxbx)
                    cell = lv[y0:y1, x0:x1]
                    ncell += 1
                    if not det_ini.detect(cell):
                        det_min.detect(cell)
        t3 = time.perf_counter()
        for lv in levels:                       # the same detector on whole levels: FAST's SIMD cost without ~1000 binding calls
            det_ini.detect(lv[16:-16, 16:-16])
        t4 = time.perf_counter()
        out.append(((t1 - t0) * 1e3, (t2 - t1) * 1e3, (t3 - t2) * 1e3, (t4 - t3) * 1e3, ncell))
    a = np.array(out)
    med = np.median(a[:, :4], axis=0)
    return {"resize_ms": float(med[0]), "blur_ms": float(med[1]), "fast_cells_ms": float(med[2]), "fast_whole_levels_ms": float(med[3]),
            "total_ms": float(np.median(a[:, :3].sum(axis=1))), "total_with_whole_level_fast_ms": float(np.median(a[:, [0, 1, 3]].sum(axis=1))),
            "fast_cells_per_frame": int(a[0, 4]), "frames": len(frames), "cv2_version": cv2.__version__,
            "note": "fast_cells_ms goes through the Python binding once per cell (an upper bound of the C++ calls); fast_whole_levels_ms is one "
                    "iniThFAST call per level (a lower bound: no minThFAST retry, no per-cell NMS borders)"}


def cpu_extract_match(frames: np.ndarray, nthreads: int, use_ref: bool, window: int | None = None, fast=False):
    """-> (seconds, kind). frames: [n,h,w] uint8 host."""
    O = oracle_mod()
    p = O.Params(PARAMS["nfeatures"], PARAMS["scale"], PARAMS["nlevels"], PARAMS["ini_th"], PARAMS["min_th"])
    L = ref_lib(fast) if use_ref else None
    if fast and L is None:
        return None, None
    frames = np.ascontiguousarray(frames)
    n, h, w = frames.shape
    tm, nc = C.c_int64(), C.c_uint64()
    fn = L.ref_extract_match_many if L is not None else O.lib().oracle_extract_match_many
    t0 = time.perf_counter()
    fn(C.byref(p), frames.ctypes.data_as(C.c_void_p), C.c_size_t(frames.strides[0]), C.c_int(w), C.c_int(h),
       C.c_size_t(frames.strides[1]), C.c_int(n), C.c_int(nthreads), C.c_int(WINDOW if window is None else window),
       C.c_float(NNRATIO), C.c_int(1),
       C.byref(tm), C.byref(nc))
    return time.perf_counter() - t0, ("reference" if L is not None else "port")


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from orb_slam2_map_b200 import synth
    cores = os.cpu_count() or 1
    sample = max(16 * cores, 64)          # 16 frames per thread and step: long enough for a steady rate, a few seconds for the default run
    hd = args.workload == "hd"
    if hd:
        sample = max(2 * cores, 8)
    win = 0 if hd else None           # window 0: no candidates, i.e. extraction only
    frames = synth.frame_sequence(sample, W, H, seed=0)
    # the build the reference's own CMakeLists asks for (-O3 -march=native, here the host's x86-64 level) when it exists for this
    # CPU; the -O2 parity build otherwise
    fast = ref_lib(fast=True) is not None
    for _ in range(args.warmup):
        cpu_extract_match(frames[:max(cores, 8)], cores, True, win, fast=fast)
    t, kind = 0.0, "port"
    for _ in range(args.steps):
        dt, kind = cpu_extract_match(frames, cores, True, win, fast=fast)
        t += dt
    fps = sample * args.steps / t
    line = {
        "impl": "reference", "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": (f"configs[3]: {W}x{H} frames, nFeatures {PARAMS['nfeatures']}, extraction only" if hd else
                                "configs[1]: 640x480 frames, nFeatures 1000, extract + consecutive-frame SearchForInitialization"),
                   "frames_per_step": sample,
                   "implementation": ("reference src/ORBextractor.cc compiled unmodified (oracle/_ref, " +
                                      (f"-O3 -march=x86-64-{cpu_level()} as CMakeLists.txt:14-15 asks" if fast else "-O2 parity build") +
                                      ") over the scalar cv models + oracle SearchForInitialization") if kind == "reference" else "oracle port"},
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": cores, "kind": kind,
                         "sample": f"{sample} frames + {sample - 1} pairs per step, one frame per thread"},
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)



# ---- brute-force 2-NN (BASELINE.json configs[4]) ---------------------------------------------------
def int_pipe_roofline(dev):
    """MEASURED integer-pipe rates of this GPU (orbm_measure_int_rates: register-only popc / lop3 / add loops and K10's
    inner loop on register operands).  The last one is the roofline of knn2_kernel: what the XOR/CSA/POPC/top-2 mix of one
    compare costs when no shared-memory load, barrier or tail is in the way."""
    import torch
    from orb_slam2_map_b200 import capi
    props = torch.cuda.get_device_properties(dev)
    r = capi.measure_int_rates()
    return {"peak": r["knn_mix"], "unit": "cmp/s", "peak_source": "measured: K10 inner loop on register operands (orbm_measure_int_rates)",
            "popc_per_s": r["popc"], "lop3_per_s": r["lop3"], "add_per_s": r["add"], "sms": props.multi_processor_count,
            "popc_per_clk_per_sm_at_1965MHz": r["popc"] / props.multi_processor_count / 1.965e9,
            "lop3_per_clk_per_sm_at_1965MHz": r["lop3"] / props.multi_processor_count / 1.965e9}


def hamming_leg(dev, rank, world, stream, queries=1 << 20, per=2000000, cpu=False):
    """The cmp/s half of BASELINE.json's metric, MEASURED on the ranks of this run: config 5 as named — 1 048 576 queries
    replicated, 2 000 000 database rows per GPU (16 M over 8) — through ONE C-ABI call per rank, orbm_knn2_sharded: local
    scan (K10), ncclAllGather of the 12-byte per-query records, device merge (K11), chunked so that the exchange of a chunk
    overlaps the scan of the next.  One timed step (2.1e12 compares per rank), CUDA events on the launching stream, max over
    ranks.  Also timed at N > 1: the exchange alone (empty shards) and the same scan without exchange (the N = 1 workload: weak
    scaling keeps the per-rank work fixed), whose ratio to the sharded per-GPU rate is the efficiency."""
    import torch
    import torch.distributed as dist
    from orb_slam2_map_b200 import sharding, synth
    q = torch.from_numpy(synth.random_descriptors(queries, 1)).to(dev)
    db = torch.from_numpy(synth.random_descriptors(per, 100 + rank)).to(dev)
    od = torch.empty((queries, 2), dtype=torch.int16, device=dev)
    oi = torch.empty((queries, 2), dtype=torch.int32, device=dev)
    comm = sharding.ShardComm(dev) if world > 1 else None
    st = stream.cuda_stream

    def timed(fn):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        fn()
        e1.record(stream)
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    small = 65536
    for _ in range(2):                                                  # warm-up: same path, 1/16 of the queries
        sharding.knn2_sharded(q[:small], db, rank * per, comm, stream=st, out=(od[:small], oi[:small]))
    ms = timed(lambda: sharding.knn2_sharded(q, db, rank * per, comm, stream=st, out=(od, oi)))
    checksum = int(oi[:, 0].to(torch.int64).sum().item())                # rank 0's view of the merged result
    ms_exchange = timed(lambda: sharding.knn2_sharded(q, db[:0], rank * per, comm, stream=st, out=(od, oi))) if world > 1 else 0.0
    # the N = 1 workload on this rank: the same scan without communicator (at N = 1 that IS the step just timed)
    ms_local = timed(lambda: sharding.knn2_sharded(q, db, rank * per, None, stream=st, out=(od, oi))) if world > 1 else ms
    if comm is not None:
        comm.close()
    value = float(queries) * per * world / (ms * 1e-3)
    local_rate = float(queries) * per / (ms_local * 1e-3)
    roof = int_pipe_roofline(dev)
    roof["achieved"] = value / world
    roof["frac"] = value / world / roof["peak"]
    cpu_sample = None
    if cpu and rank == 0 and world == 1:
        # config 5 on the host cores: the oracle's brute-force top-2 (__builtin_popcountll, all threads) on a bounded sample of the
        # same descriptors, its own result checked against the GPU's for those queries
        cores = os.cpu_count() or 1
        nq_c, ndb_c = 2048, min(per, 1000000)
        O = oracle_mod()
        qc, dbc = q[:nq_c].cpu().numpy(), db[:ndb_c].cpu().numpy()
        O.knn2(qc[:64], dbc[:65536], 0, cores)
        t0 = time.perf_counter()
        od_c, oi_c = O.knn2(qc, dbc, 0, cores)
        dt_c = time.perf_counter() - t0
        god = torch.empty((nq_c, 2), dtype=torch.int16, device=dev); goi = torch.empty((nq_c, 2), dtype=torch.int32, device=dev)
        sharding.knn2_sharded(q[:nq_c], db[:ndb_c], 0, None, stream=st, out=(god, goi))
        torch.cuda.synchronize()
        cpu_sample = {"value": float(nq_c) * ndb_c / dt_c, "unit": "cmp/s", "cores": cores, "kind": "port",
                      "sample": f"{nq_c} queries x {ndb_c} rows of the same data, oracle_knn2 on {cores} threads",
                      "equals_gpu_result": bool(np.array_equal(goi.cpu().numpy(), oi_c) and np.array_equal(god.cpu().numpy().view(np.uint16), od_c))}
    return {"metric": "Hamming cmp/s (brute-force top-2 kNN, 256-bit descriptors)", "value": value, "unit": "cmp/s", "per_gpu": value / world,
            "cpu_sample": cpu_sample,
            "measured": True, "nranks": world, "queries": queries, "database_rows_per_gpu": per, "database_rows_total": per * world,
            "steps": 1, "ms": ms, "exchange_ms_alone": ms_exchange, "local_scan_cmp_s_per_gpu": local_rate,
            "efficiency_vs_n1": value / world / local_rate,
            "how": "orbm_knn2_sharded (C ABI): K10 scan of the rank's shard, ncclAllGather of 12 B per query and rank, K11 merge; "
                   "2 query chunks, the exchange of a chunk overlaps the scan of the next; max over ranks of CUDA-event time",
            "merged_top1_index_checksum": checksum, "popc_roofline": roof}


def run_knn(args):
    """configs[4]: queries replicated, database rows sharded over the ranks (weak scaling: rows per GPU fixed),
    local top-2 on every rank, all_gather of 12 B per query and rank over NCCL, device merge."""
    import torch
    import torch.distributed as dist
    from orb_slam2_map_b200 import sharding, synth
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    numa = bind_to_gpu_numa(local) if world > 1 else {"bound": False}
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev) if world > 1 else None
    if world == 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29533")
        dist.init_process_group("nccl", rank=0, world_size=1, device_id=dev)
    nq, per = args.queries, args.db_per_gpu
    q = torch.from_numpy(synth.random_descriptors(nq, 1)).to(dev)
    db = torch.from_numpy(synth.random_descriptors(per, 100 + rank)).to(dev)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    comm = sharding.ShardComm(dev) if world > 1 else None
    step = lambda: sharding.knn2_sharded(q, db, rank * per, comm, stream=stream.cuda_stream)
    for _ in range(max(args.warmup, 1)):
        step()
    clocks = ClockSampler(local)
    dist.barrier(); torch.cuda.synchronize()
    clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dist.barrier(); torch.cuda.synchronize()
    e0.record(stream)
    for _ in range(args.steps):
        od, oi = step()
    e1.record(stream)
    dist.barrier(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    clk = clocks.stop()
    # e2e: queries from pinned host memory in, merged result to the host out
    hq = torch.empty((nq, 32), dtype=torch.uint8, pin_memory=True); hq.copy_(q)
    hod = torch.empty((nq, 2), dtype=torch.int16, pin_memory=True); hoi = torch.empty((nq, 2), dtype=torch.int32, pin_memory=True)
    def e2e_step():
        q.copy_(hq, non_blocking=True)
        d, i = step()
        hod.copy_(d, non_blocking=True); hoi.copy_(i, non_blocking=True)
        stream.synchronize()
    e2e_step(); dist.barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    total = float(nq) * per * world
    value = total * args.steps / (ms * 1e-3)
    roof = int_pipe_roofline(dev)
    peak = roof["peak"]
    line = {"metric": "Hamming cmp/s (brute-force top-2 kNN, 256-bit descriptors)", "value": value, "unit": "cmp/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32 xor/popc", "data": "synthetic",
            "config": {"workload": "configs[4]: brute-force Hamming top-2, queries replicated, database sharded by rows, NCCL all_gather of "
                                   "per-query top-2 + device merge", "queries": nq, "database_rows_per_gpu": per,
                       "l2_policy": f"database shard {per * 32 / 1e6:.0f} MB streamed once per query tile"},
            "clocks": clk,
            "e2e": {"value": total * args.steps / float(dt.item()), "unit": "cmp/s", "h2d_bytes_per_step": nq * 32, "d2h_bytes_per_step": nq * 12},
            "gpu_launches": 2 * args.steps,
            "roofline": {"bound": "int-pipe", "kernel": "knn2_kernel", "achieved": value / world, "peak": peak, "unit": "cmp/s",
                         "frac": value / world / peak, "traffic": None,
                         "how": roof["peak_source"], "int_rates": roof}}
    if rank == 0:
        emit(line)
    dist.destroy_process_group()


def run_stereo(args):
    """configs[2]: KITTI-shaped 1241x376 stereo pairs, nFeatures 2000: left+right extraction + ComputeStereoMatches
    with the pyramids kept on the device; pairs sharded by rank, no collective."""
    import torch
    import torch.distributed as dist
    from orb_slam2_map_b200 import capi, synth
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    numa = bind_to_gpu_numa(local) if world > 1 else {"bound": False}
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    w, h, P = 1241, 376, (args.frames if args.frames != 4096 else 1024)
    bf, fx = 386.1448, 718.856
    base = [synth.stereo_pair(w, h, seed=1000 * rank + i) for i in range(16)]
    hl = torch.empty((P, h, w), dtype=torch.uint8, pin_memory=True); hr = torch.empty((P, h, w), dtype=torch.uint8, pin_memory=True)
    for i in range(P):
        hl[i] = torch.from_numpy(base[i % 16][0]); hr[i] = torch.from_numpy(base[i % 16][1])
    dl, dr = hl.to(dev), hr.to(dev)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    ex = capi.Extractor(nfeatures=2000, max_width=w, max_height=h, chunk_frames=min(args.chunk, 512), result_slots=2 * P, device=local,
                        stream=stream.cuda_stream)
    step = lambda: ex.extract_stereo_batch(dl.data_ptr(), dr.data_ptr(), bf, fx, nframes=P, width=w, height=h, frame_stride=w * h,
                                           row_stride=w, on_device=True, fetch=False)
    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    for _ in range(args.warmup):
        step()
    ex.sync(); ex.profile(); ex.set_profiling(True)
    clocks = ClockSampler(local)
    barrier(); clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    barrier()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    stage_ms, stage_launches = ex.profile(); ex.set_profiling(False)
    clk = clocks.stop()
    hln, hrn = hl.numpy(), hr.numpy()
    outs = (torch.empty((P, ex.cap), dtype=torch.float32, pin_memory=True).numpy(),
            torch.empty((P, ex.cap), dtype=torch.float32, pin_memory=True).numpy())        # mvuRight, mvDepth in pinned host memory
    ex.extract_stereo_batch(hln, hrn, bf, fx, out=outs)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ur, dp, ncmp = ex.extract_stereo_batch(hln, hrn, bf, fx, out=outs)
    dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    value = world * P * args.steps / (ms * 1e-3)
    peak, peak_src = measured_peaks()
    ab = algorithmic_bytes(w, h, 2005)["total"] * 2
    line = {"metric": "stereo pairs/s (1241x376, 2000 kp, L+R extract + ComputeStereoMatches)", "value": value, "unit": "pairs/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "configs[2]: KITTI-shaped 1241x376 stereo pairs, nFeatures 2000, left/right extraction + ComputeStereoMatches",
                       "pairs_per_step_per_gpu": P, "l2_policy": f"inputs ({2 * P * w * h / 1e9:.2f} GB per step) larger than L2",
                       "sharding": "pairs by rank, no collective", "rank_cpu_binding": numa},
            "clocks": clk,
            "e2e": {"value": world * P * args.steps / float(dt.item()), "unit": "pairs/s", "h2d_bytes_per_step": 2 * P * w * h,
                    "d2h_bytes_per_step": 2 * P * ex.cap * 4 + 8},
            "gpu_launches": int(sum(stage_launches.values())),
            "roofline": {"bound": "hbm", "kernel": "path", "achieved": ab * value / world / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": ab * value / world / 1e9 / peak, "traffic": None, "peak_source": peak_src,
                         "algorithmic_bytes_per_pair": ab},
            "stages": {k: {"ms_per_step": v / args.steps, "launches_per_step": stage_launches[k] / args.steps}
                       for k, v in stage_ms.items() if stage_launches[k]},
            "hamming_compares_per_step": int(ncmp), "stereo_matches_per_pair": float((ur >= 0).sum() / P)}
    if rank == 0:
        emit(line)
    ex.close()
    if world > 1:
        dist.destroy_process_group()


def run_rgbd(args):
    """Side line for SURVEY 8(f) rank 1: the RGB-D frame front (colour -> gray, ORB, mvKeysUn, mvuRight/mvDepth, grid)
    on TUM-shaped input: 640x480 RGB + 16-bit depth, TUM1 calibration (Examples/RGB-D/TUM1.yaml)."""
    import torch
    from orb_slam2_map_b200 import capi, synth
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    F = args.frames if args.frames != 4096 else 2048
    gray = synth.frame_sequence_torch(F, W, H, seed=0, device=dev)
    g = torch.Generator(device=dev); g.manual_seed(1)
    color = torch.stack([(gray.to(torch.int16) + torch.randint(-12, 13, gray.shape, generator=g, device=dev, dtype=torch.int16)).clamp(0, 255).to(torch.uint8)
                         for _ in range(3)], dim=-1).contiguous()
    depth = torch.randint(500, 30000, (F, H, W), generator=g, device=dev, dtype=torch.int32).to(torch.uint16)
    cam = capi.Camera(517.306408, 516.469215, 318.643040, 255.313989, 0.262383, -0.953104, -0.005358, 0.002628, 1.163314, 40.0)
    factor = 1.0 / 5000.0
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    ex = capi.Extractor(**PARAMS, max_width=W, max_height=H, chunk_frames=min(args.chunk, 512), result_slots=F, device=0, stream=stream.cuda_stream)
    step = lambda: ex.frames_rgbd_ptr(color.data_ptr(), 3, depth.data_ptr(), True, F, W, H, cam, True, factor, True, 0)
    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    clocks = ClockSampler(0); clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    clk = clocks.stop()
    hc = torch.empty(color.shape, dtype=torch.uint8, pin_memory=True); hc.copy_(color)
    hd = torch.empty(depth.shape, dtype=torch.uint16, pin_memory=True); hd.copy_(depth)
    torch.cuda.synchronize()
    # results land in page-locked buffers that are allocated once (a caller's Frame objects own theirs)
    pin = lambda shape, dt: torch.empty(shape, dtype=dt, pin_memory=True).numpy()
    cap = ex.cap
    o_kp = pin((F, cap, 28), torch.uint8).view(capi.KP_DTYPE).reshape(F, cap); o_de = pin((F, cap, 32), torch.uint8)
    o_un = pin((F, cap, 28), torch.uint8).view(capi.KP_DTYPE).reshape(F, cap); o_ur = pin((F, cap), torch.float32); o_z = pin((F, cap), torch.float32)
    def e2e():
        ex.frames_rgbd_ptr(hc.data_ptr(), 3, hd.data_ptr(), True, F, W, H, cam, True, factor, False, 0)
        ex.download(0, F, out=(o_kp, o_de)); ex.download_frame_front(0, F, out=(o_un, o_ur, o_z))
    e2e()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e()
    dt = time.perf_counter() - t0
    peak, peak_src = measured_peaks()
    ab = algorithmic_bytes(W, H, 1005)["total"] + 4 * W * H          # + colour read (3 B/px) and gray write (1 B/px)
    value = F * args.steps / (ms * 1e-3)
    line = {"metric": "RGB-D frames/s (640x480 RGB + 16-bit depth, 1000 kp: cvtColor + ORB + UndistortKeyPoints + ComputeStereoFromRGBD + grid)",
            "value": value, "unit": "frames/s", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "SURVEY 8(f) rank 1: RGB-D frame front on TUM1-shaped input", "frames_per_step_per_gpu": F,
                       "chunk_frames": min(args.chunk, 512), "l2_policy": f"inputs ({F * W * H * 5 / 1e9:.2f} GB per step) larger than L2"},
            "clocks": clk,
            "e2e": {"value": F * args.steps / dt, "unit": "frames/s", "h2d_bytes_per_step": F * W * H * 5,
                    "d2h_bytes_per_step": F * ex.cap * (28 + 32 + 28 + 8) + 4 * F},
            # per chunk: cvt_gray, 7 resize, blur, FAST, quadtree, describe, depth convertTo, frame_front (+ grid build)
            "gpu_launches": args.steps * (-(-F // min(args.chunk, 512))) * (PARAMS["nlevels"] - 1 + 8),
            "roofline": {"bound": "hbm", "kernel": "path", "achieved": ab * value / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": ab * value / 1e9 / peak, "traffic": None, "peak_source": peak_src, "algorithmic_bytes_per_frame": ab}}
    emit(line)
    ex.close()


def run_rgbd_single(args):
    """The fork's main mode, one frame per call: Tracking::GrabImageRGBD + Frame::Frame(RGB-D) (src/Tracking.cc:300-339,
    src/Frame.cc:136-189) through orbx_frame_rgbd — pageable RGB + 16-bit depth in, mvKeys / mvKeysUn / descriptors / mvuRight /
    mvDepth out.  CPU side, one thread: the oracle's cvtColor + the reference's own ORBextractor.cc (oracle/_ref) +
    undistortPoints + ComputeStereoFromRGBD."""
    import torch
    from orb_slam2_map_b200 import capi, synth
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local)
    NF = 32
    gray = synth.frame_sequence(NF, W, H, seed=0)
    rng = np.random.default_rng(3)
    # R = G = B = the synthetic frame: cvtColor returns exactly that frame, so the extractor sees the images of the `single` line
    color = np.ascontiguousarray(np.repeat(gray[..., None], 3, axis=-1))
    depth = rng.integers(0, 30000, (NF, H, W), dtype=np.uint16)
    factor = float(np.float32(1.0) / np.float32(5000.0))
    tum1 = dict(fx=517.306408, fy=516.469215, cx=318.643040, cy=255.313989, k1=0.262383, k2=-0.953104, p1=-0.005358, p2=0.002628, k3=1.163314, bf=40.0)
    cam = capi.Camera(**tum1)
    ex = capi.Extractor(**PARAMS, max_width=W, max_height=H, chunk_frames=1, result_slots=1, device=local)
    L = capi.lib()
    cap = ex.reserve_shape(W, H)
    kps = np.zeros(cap, capi.KP_DTYPE); kun = np.zeros(cap, capi.KP_DTYPE); desc = np.zeros((cap, 32), np.uint8)
    ur = np.zeros(cap, np.float32); z = np.zeros(cap, np.float32); b = np.zeros(4, np.float32); n = C.c_int(0)
    P = lambda a: a.ctypes.data_as(C.c_void_p)

    # argument objects built once: the call below is what a C++ caller pays, not ctypes marshalling of 20 arguments
    L.orbx_frame_rgbd.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_void_p, C.c_int, C.c_float, C.c_size_t, C.c_int,
                                  C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
    outs = (kps.ctypes.data, kun.ctypes.data, desc.ctypes.data, ur.ctypes.data, z.ctypes.data)
    camp, np_, bp = C.addressof(cam), C.addressof(n), b.ctypes.data

    def call(c, d):
        rc = L.orbx_frame_rgbd(ex.h, c.ctypes.data, 3, 1, W * 3, d.ctypes.data, 1, factor, W * 2, W, H, camp, outs[0], outs[1], outs[2], outs[3],
                               outs[4], cap, np_, bp)
        if rc:
            capi.check(rc)

    def measure(cs, ds, calls):
        for i in range(40):
            call(cs[i % NF], ds[i % NF])
        lat = np.empty(calls)
        for i in range(calls):
            t0 = time.perf_counter()
            call(cs[i % NF], ds[i % NF])
            lat[i] = time.perf_counter() - t0
        return lat

    calls = max(args.steps, 1) * 200
    clocks = ClockSampler(local); clocks.start()
    t_all = time.perf_counter()
    lat = measure(color, depth, calls)
    t_all = time.perf_counter() - t_all
    clk = clocks.stop()
    pc = torch.empty(color.shape, dtype=torch.uint8, pin_memory=True); pc.copy_(torch.from_numpy(color))
    pd = torch.empty(depth.shape, dtype=torch.uint16, pin_memory=True); pd.copy_(torch.from_numpy(depth))
    lat_p = measure(pc.numpy(), pd.numpy(), calls)
    # the batch entry point called with one frame (the round-1 way): plain launches, three downloads
    def old(i):
        ex.frames_rgbd(color[i % NF:i % NF + 1], depth[i % NF:i % NF + 1], cam, rgb_order=True, depth_factor=factor)
        ex.download(0, 1); ex.download_frame_front(0, 1)
    for i in range(10):
        old(i)
    lat_o = np.empty(100)
    for i in range(100):
        t0 = time.perf_counter(); old(i); lat_o[i] = time.perf_counter() - t0
    med = float(np.median(lat)) * 1e6
    line = {"metric": "RGB-D single-frame latency (640x480 RGB + 16-bit depth, 1000 kp: GrabImageRGBD + Frame front)", "value": med, "unit": "us/frame",
            "n_gpus": 1, "steps": args.steps, "warmup": args.warmup, "ms_per_step": t_all / (calls + 40) * 1e3, "higher_is_better": False,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "one RGB-D frame per call through orbx_frame_rgbd: pageable colour + depth in, mvKeys, mvKeysUn, descriptors, "
                                   "mvuRight, mvDepth out (CUDA-graph replay of the extractor, one synchronisation)",
                       "calls": calls, "distinct_frames": NF, "keypoints": int(n.value), "l2_policy": "latency line: one frame fits L2 by definition"},
            "clocks": clk,
            "latency_us": {"median": med, "p10": float(np.percentile(lat, 10)) * 1e6, "p90": float(np.percentile(lat, 90)) * 1e6,
                           "p99": float(np.percentile(lat, 99)) * 1e6, "pinned_source_median": float(np.median(lat_p)) * 1e6,
                           "batch_entry_point_one_frame_median": float(np.median(lat_o)) * 1e6},
            "e2e": {"value": med, "unit": "us/frame", "h2d_bytes_per_step": W * H * 5, "d2h_bytes_per_step": cap * (28 + 28 + 32 + 8) + 8},
            "gpu_launches": (calls + 40) * (1 + PARAMS["nlevels"] - 1 + 3 * PARAMS["nlevels"] + 1 + 2)}
    if not args.no_cpu:
        O = oracle_mod()
        camv = np.array([tum1[k] for k in ("fx", "fy", "cx", "cy")], np.float32)
        dist = np.array([tum1[k] for k in ("k1", "k2", "p1", "p2", "k3")], np.float32)
        ts = []
        for i in range(24):
            t0 = time.perf_counter()
            g = O.cvt_gray(color[i % NF], True)
            t1 = time.perf_counter()
            ext = cpu_single_thread_ms(g[None], fast=False)[0]
            t2 = time.perf_counter()
            ko = kps[:int(n.value)]
            xy = np.stack([ko["x"], ko["y"]], axis=1).astype(np.float32)
            xyu = O.undistort_points(xy, camv, dist)
            dflt = O.depth_to_float(depth[i % NF], factor)
            O.stereo_from_rgbd(xy, xyu, dflt, tum1["bf"])
            t3 = time.perf_counter()
            ts.append(((t1 - t0) * 1e3 + ext + (t3 - t2) * 1e3, ext))
        a = np.array(ts)
        line["cpu_baseline"] = {"value": float(np.median(a[:, 0])) * 1e3, "unit": "us/frame", "cores": 1, "kind": "reference",
                                "sample": "24 frames, one thread: oracle cvtColor + the reference's ORBextractor.cc (oracle/_ref, -O2 parity build: "
                                          f"{float(np.median(a[:, 1])):.2f} ms of it) + oracle undistortPoints / convertTo / ComputeStereoFromRGBD"}
    emit(line)
    ex.close()


def run_matcher_calls(args):
    """The stateless per-frame / per-keyframe matcher calls of include/orbx.h as Tracking, LocalMapping and LoopClosing make
    them: host structs in, host results out, ONE frame or keyframe pair per call.  Each call's median latency next to the
    CPU oracle's on the same inputs (single thread, as ORB-SLAM2 runs them).  Inputs: the case builders of tests/."""
    import torch
    from orb_slam2_map_b200 import capi, synth
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    O = oracle_mod()          # input containers of the case builders + the CPU leg; the frames below come from the GPU extractor
    import test_match2_gpu as T
    from test_oracle_kfmatch import make_fuse_case, make_sim3_case, make_tri_pair
    from test_oracle_bow import make_pair
    from test_kfmatch_gpu import _to_capi

    def med(fn, n, warm=5):
        for _ in range(warm):
            fn()
        ts = np.empty(n)
        for i in range(n):
            t0 = time.perf_counter(); fn(); ts[i] = time.perf_counter() - t0
        return float(np.median(ts)) * 1e6

    calls = max(args.steps, 1) * 40
    cpu_calls = 12
    rows = {}

    def add(name, gpu_fn, cpu_fn, **info):
        g = med(gpu_fn, calls)
        row = {"gpu_us": round(g, 1), **info}
        if not args.no_cpu:
            c = med(cpu_fn, cpu_calls, warm=2)
            row.update(cpu_oracle_us=round(c, 1), speedup=round(c / g, 2))
        rows[name] = row

    ex = capi.Extractor(**PARAMS, max_width=640, max_height=480, chunk_frames=1, result_slots=1, device=local)
    sf = ex.tables()["scale"]
    k, d = ex.extract(synth.textured_canvas(640, 480, seed=1))
    k, d, q, qd, ur, occ, blk, sf = T._queries_from_keypoints(k, d, sf, 1)
    fo = O.FrameData(k, d, 640, 480, u_right=ur, scale_factors=sf)
    fg = capi.FrameData(k, d, 640, 480, u_right=ur, scale_factors=sf)
    for name, variant, th, ratio, ori in (("search_by_projection_localmap", capi.PROJ_LOCALMAP, 100, 0.8, False),
                                          ("search_by_projection_lastframe", capi.PROJ_LASTFRAME, 100, 0.9, True),
                                          ("search_by_projection_keyframe", capi.PROJ_KEYFRAME, 64, 0.9, True),
                                          ("search_by_projection_sim3", capi.PROJ_SIM3, 50, 0.75, False)):
        qq = q
        if variant in (capi.PROJ_KEYFRAME, capi.PROJ_SIM3):
            qq = q.copy(); qq["min_level"] = -1; qq["max_level"] = -1
            qq["level_lo"] = np.maximum(k["octave"][:len(q)] - 1, 0); qq["level_hi"] = qq["level_lo"] + 1
        add(name, lambda: capi.search_by_projection(variant, fg, qq, qd, blk, occ, th, ratio, ori),
            lambda: O.search_by_projection(variant, fo, qq, qd, blk, occ, th, ratio, ori), queries=len(q), keypoints=len(k))
    frames = synth.frame_sequence(2, 640, 480, seed=3)
    ka, da = ex.extract(frames[0]); kb, db = ex.extract(frames[1])
    prev = np.stack([ka["x"], ka["y"]], axis=1)
    f1o, f2o = O.FrameData(ka, da, 640, 480), O.FrameData(kb, db, 640, 480)
    f1g, f2g = capi.FrameData(ka, da, 640, 480), capi.FrameData(kb, db, 640, 480)
    add("search_for_initialization", lambda: capi.search_for_initialization(f1g, f2g, prev, 100, 0.9, True),
        lambda: O.search_for_initialization(f1o, f2o, prev, 100, 0.9, True), keypoints=[len(ka), len(kb)])
    left, right = synth.stereo_pair(1241, 376, seed=5)
    sp_params = dict(PARAMS, nfeatures=2000)
    exs = capi.Extractor(**sp_params, max_width=1241, max_height=376, chunk_frames=1, result_slots=1, device=local)
    kl, dl = exs.extract(left); pl = [exs.level(0, l) for l in range(8)]
    kr, dr = exs.extract(right); pr = [exs.level(0, l) for l in range(8)]
    tb = exs.tables()
    st = (kl, dl, kr, dr, pl, pr, tb["scale"], tb["inv_scale"], 386.1448, 718.856)
    add("compute_stereo_matches_kitti", lambda: capi.compute_stereo_matches(*st), lambda: O.compute_stereo_matches(*st),
        keypoints=[len(kl), len(kr)])
    exs.close()
    voc = O.make_vocabulary(k=10, depth=4, seed=3)
    A, B = make_pair(voc, n=1000, seed=1, levelsup=2)
    for v in (0, 1):
        add(f"search_by_bow_{'frame' if v == 0 else 'keyframes'}", lambda: capi.search_by_bow(v, [(A, B)], nnratio=0.7, check_ori=True),
            lambda: O.search_by_bow(v, *A, *B, nnratio=0.7, check_ori=True), features=1000)
    kf, fq, fqd = make_fuse_case(25, n=2000, nq=3000)
    kfg = _to_capi(kf)
    add("fuse", lambda: capi.fuse(0, kfg, fq, fqd), lambda: O.fuse(0, kf, fq, fqd), keypoints=2000, queries=3000)
    s3 = make_sim3_case(21)
    s3g = (_to_capi(s3[0]), _to_capi(s3[1])) + tuple(s3[2:])
    add("search_by_sim3", lambda: capi.search_by_sim3(*s3g), lambda: O.search_by_sim3(*s3), queries=[len(s3[2]), len(s3[5])])
    ta, tb, F12, epi = make_tri_pair(4, n1=2000, n2=1900, mono_frac=0.0)
    tg = [(_to_capi(ta), _to_capi(tb))]
    add("search_for_triangulation", lambda: capi.search_for_triangulation(tg, [F12], [epi], False, True),
        lambda: O.search_for_triangulation(ta, tb, F12, epi, False, True), keypoints=[2000, 1900])
    rng = np.random.default_rng(8)
    sizes = rng.integers(2, 40, 300)
    offs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    dd = rng.integers(0, 256, (int(offs[-1]), 32), dtype=np.uint8)
    add("distinctive_descriptors_300_points", lambda: capi.distinctive_descriptors(dd, offs), lambda: O.distinctive_descriptors(dd, offs),
        observations=int(offs[-1]))
    ex.close()
    sp = [r["speedup"] for r in rows.values() if "speedup" in r]
    head = rows["search_by_projection_localmap"]
    line = {"metric": "stateless matcher call latency (SearchByProjection local map, 900 map points x 1008 keypoints)", "value": head["gpu_us"],
            "unit": "us/call", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup, "ms_per_step": head["gpu_us"] / 1e3,
            "higher_is_better": False, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "stateless orbm_* calls, one frame / keyframe pair per call, host structs in and out (includes the ctypes "
                                   "marshalling of orb_slam2_map_b200/capi.py on the GPU side and of tests/oracle_lib.py on the CPU side)",
                       "calls_per_entry": calls, "l2_policy": "latency line: one frame fits L2 by definition"},
            "calls": rows, "min_speedup_vs_cpu_oracle": min(sp) if sp else None,
            "e2e": {"value": head["gpu_us"], "unit": "us/call", "h2d_bytes_per_step": int(k.nbytes + d.nbytes + q.nbytes + qd.nbytes + len(k) * 5 + len(q)),
                    "d2h_bytes_per_step": int(len(k) * 5 + len(q) * 8 + 4)},
            "gpu_launches": calls * 2}
    if not args.no_cpu:
        line["cpu_baseline"] = {"value": head.get("cpu_oracle_us"), "unit": "us/call", "cores": 1, "kind": "port",
                                "sample": f"{cpu_calls} calls per entry, oracle/ through tests/oracle_lib.py"}
    emit(line)


def run_single(args):
    """BASELINE configs[0] / SURVEY 8(d) "Config 1": ONE 640x480 frame per call, the way Tracking uses the extractor
    (Frame::ExtractORB, src/Frame.cc:247-253): pageable host image in, host keypoints + descriptors out, through
    orbx_extract.  The library replays a CUDA graph of the frame's kernels; the number is the latency the caller of
    ORBextractor::operator() sees.  CPU side: the reference's own ORBextractor.cc (oracle/_ref), single-threaded per frame."""
    import torch
    from orb_slam2_map_b200 import capi, synth
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    NF = 64
    frames = synth.frame_sequence_torch(NF, W, H, seed=rank, device=dev).cpu().numpy()     # pageable host memory, like a cv::Mat
    ex = capi.Extractor(**PARAMS, max_width=W, max_height=H, chunk_frames=1, result_slots=1, device=local)  # own stream -> graph path
    L = capi.lib()
    kps = np.zeros(ex.cap, capi.KP_DTYPE); desc = np.zeros((ex.cap, 32), np.uint8); n = C.c_int(0)
    pk, pd = kps.ctypes.data_as(C.c_void_p), desc.ctypes.data_as(C.c_void_p)

    def one(i):
        f = frames[i % NF]
        capi.check(L.orbx_extract(ex.h, f.ctypes.data_as(C.c_void_p), W, H, C.c_size_t(W), pk, pd, ex.cap, C.byref(n)))

    calls = max(args.steps, 1) * 200
    for i in range(max(args.warmup, 3) * 20):
        one(i)
    clocks = ClockSampler(local); clocks.start()
    lat = np.empty(calls)
    t_all = time.perf_counter()
    for i in range(calls):
        t0 = time.perf_counter()
        one(i)
        lat[i] = time.perf_counter() - t0
    t_all = time.perf_counter() - t_all
    clk = clocks.stop()
    nkp = int(n.value)
    # the same calls with the images in page-locked host memory (no staging copy inside the library)
    pinned = torch.empty((NF, H, W), dtype=torch.uint8, pin_memory=True)
    pinned.copy_(torch.from_numpy(frames))
    pnp = pinned.numpy()

    def one_pinned(i):
        f = pnp[i % NF]
        capi.check(L.orbx_extract(ex.h, f.ctypes.data_as(C.c_void_p), W, H, C.c_size_t(W), pk, pd, ex.cap, C.byref(n)))

    for i in range(40):
        one_pinned(i)
    lat_p = np.empty(calls)
    for i in range(calls):
        t0 = time.perf_counter()
        one_pinned(i)
        lat_p[i] = time.perf_counter() - t0
    # the same frame's kernels alone (device-resident image, no graph: 12 launches), CUDA events on the launching stream
    stream = torch.cuda.Stream(device=dev); torch.cuda.set_stream(stream)
    ex2 = capi.Extractor(**PARAMS, max_width=W, max_height=H, chunk_frames=1, result_slots=1, device=local, stream=stream.cuda_stream)
    dfr = torch.from_numpy(frames[:8]).to(dev)
    for i in range(20):
        ex2.extract_batch(dfr[i % 8].data_ptr(), nframes=1, width=W, height=H, frame_stride=W * H, row_stride=W, on_device=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for i in range(200):
        ex2.extract_batch(dfr[i % 8].data_ptr(), nframes=1, width=W, height=H, frame_stride=W * H, row_stride=W, on_device=True)
    e1.record(stream)
    torch.cuda.synchronize()
    dev_us = e0.elapsed_time(e1) * 1e3 / 200
    med = float(np.median(lat)) * 1e6
    nl = PARAMS["nlevels"]
    line = {"metric": "ORB single-frame latency (640x480, 1000 kp, ORBextractor::operator())", "value": med, "unit": "us/frame",
            "n_gpus": 1, "steps": args.steps, "warmup": args.warmup, "ms_per_step": t_all / calls * 1e3, "higher_is_better": False,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "configs[0]: single synthetic 640x480 frame per call, ORBextractor(1000,1.2,8,20,7), pageable host image in, "
                                   "host keypoints + descriptors out (orbx_extract, CUDA-graph replay)",
                       "calls": calls, "distinct_frames": NF, "keypoints": nkp, "l2_policy": "latency line: one frame fits L2 by definition"},
            "clocks": clk,
            "latency_us": {"median": med, "p10": float(np.percentile(lat, 10)) * 1e6, "p90": float(np.percentile(lat, 90)) * 1e6,
                           "p99": float(np.percentile(lat, 99)) * 1e6, "pinned_source_median": float(np.median(lat_p)) * 1e6,
                           "kernels_only_no_graph": dev_us},
            "e2e": {"value": med, "unit": "us/frame", "h2d_bytes_per_step": W * H, "d2h_bytes_per_step": ex.cap * 60 + 8,
                    "frames_per_s_one_caller": calls / t_all},
            "gpu_launches": calls * (nl - 1 + 3 * nl + 1)}
    if rank == 0 and not args.no_cpu:
        s1 = 24
        t1, kind = cpu_extract_match(frames[:s1], 1, True, 0)
        line["cpu_baseline"] = {"value": t1 / s1 * 1e6, "unit": "us/frame", "cores": 1, "kind": kind,
                                "sample": f"{s1} frames, single-threaded per frame as ORB-SLAM2 runs it (extraction only)"}
    if rank == 0:
        emit(line)
    ex.close(); ex2.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--frames", type=int, default=4096)
    ap.add_argument("--chunk", type=int, default=2048, help="frames per device-resident chunk (handle capacity); the pipeline from host images runs on 256-frame chunks inside the library (ORBX_HOST_CHUNK)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-hamming", action="store_true", help="skip the config-5 leg of the default line (profiling runs)")
    ap.add_argument("--workload", default="frames", choices=["frames", "hd", "stereo", "knn", "rgbd", "single", "rgbd_single", "matcher_calls"],
                    help="frames = BASELINE configs[1] (the graded line); hd = configs[3]; stereo = configs[2]; knn = configs[4]")
    ap.add_argument("--queries", type=int, default=262144, help="knn: queries per step (config 5 names 1M)")
    ap.add_argument("--db-per-gpu", type=int, default=2000000, help="knn: database rows per GPU (config 5: 16M over 8)")
    args = ap.parse_args()
    quiet_stdout()
    global W, H, PARAMS, METRIC
    match = True
    workload_name = ("configs[1]: batch of 4096 synthetic 640x480 frames, ORBextractor(1000,1.2,8,20,7) + "
                     "consecutive-frame SearchForInitialization(window 100, nnratio 0.9)")
    if args.workload == "hd":
        W, H = 1920, 1080
        PARAMS = dict(PARAMS, nfeatures=5000)
        METRIC = "ORB frames/s (1920x1080, 5000 kp, extract)"
        match = False
        if args.frames == 4096:
            args.frames = 512
        args.chunk = min(args.chunk, 128)
        workload_name = "configs[3]: 1920x1080 frames, ORBextractor(5000,1.2,8,20,7), extraction only"
    if args.impl == "reference":
        run_reference(args)
        return
    if args.workload == "stereo":
        run_stereo(args)
        return
    if args.workload == "knn":
        run_knn(args)
        return
    if args.workload == "rgbd":
        run_rgbd(args)
        return
    if args.workload == "single":
        run_single(args)
        return
    if args.workload == "rgbd_single":
        run_rgbd_single(args)
        return
    if args.workload == "matcher_calls":
        run_matcher_calls(args)
        return

    import torch
    import torch.distributed as dist
    from orb_slam2_map_b200 import capi, synth

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local)
    numa = bind_to_gpu_numa(local) if world > 1 else {"bound": False}
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    F = args.frames
    frames = synth.frame_sequence_torch(F, W, H, seed=rank, device=dev)
    # a real (non-default) stream: liborbx launches on it and the CUDA events below are recorded on it
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    ex = capi.Extractor(**PARAMS, max_width=W, max_height=H, chunk_frames=args.chunk, result_slots=F, device=local,
                        stream=stream.cuda_stream)

    def step_device():
        ex.extract_batch(frames.data_ptr(), nframes=F, width=W, height=H, frame_stride=W * H, row_stride=W,
                         on_device=True, first_slot=0)
        if match:
            ex.search_init_pairs(0, F - 1, WINDOW, NNRATIO, True, fetch=False)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step_device()
    ex.sync()
    ex.profile()
    ex.set_profiling(True)
    clocks = ClockSampler(local)
    barrier()
    clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for _ in range(args.steps):
        step_device()
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    stage_ms, stage_launches = ex.profile()
    ex.set_profiling(False)
    clk = clocks.stop()
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    value = world * F * args.steps / (ms * 1e-3)

    # mean keypoints per frame (for the algorithmic byte count)
    _, _, counts = ex.download(0, min(F, 64), want_kps=False, want_desc=False)
    nkp = float(counts.mean())

    # ---- e2e: host pinned images in, host results out, through the C ABI ------------------------
    host = torch.empty((F, H, W), dtype=torch.uint8, pin_memory=True)
    host.copy_(frames)
    torch.cuda.synchronize()
    cap = ex.cap
    kp_out = torch.empty((F, cap, 28), dtype=torch.uint8, pin_memory=True).numpy().view(capi.KP_DTYPE).reshape(F, cap)
    de_out = torch.empty((F, cap, 32), dtype=torch.uint8, pin_memory=True).numpy()
    m12_out = torch.empty((F - 1, cap), dtype=torch.int32, pin_memory=True).numpy()
    nm_out = torch.empty((F - 1,), dtype=torch.int32, pin_memory=True).numpy()
    hnp = host.numpy()
    ncmp = C.c_uint64()

    cnt_out = torch.empty((F,), dtype=torch.int32, pin_memory=True).numpy()

    def step_e2e():
        # host images in -> keypoints / descriptors / counts in host memory (copies overlapped with the kernels
        # chunk by chunk inside the library), the matcher of each chunk's pairs and its result copy in the same pipeline
        if match:
            # one pipelined call: the pairs of a chunk are matched right after its extraction
            ex.extract_match_to_host(hnp, kp_out, de_out, cnt_out, m12_out, nm_out, WINDOW, NNRATIO, True)
        else:
            ex.extract_batch_to_host(hnp, kp_out, de_out, cnt_out)
        return kp_out, de_out, cnt_out

    _, _, cnts = step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * F * args.steps / float(t.item())
    # the same call on PAGEABLE buffers (what a std::vector<cv::Mat> and the Frame members are): the driver stages every copy through
    # its own page-locked buffers, so this is the rate a caller sees without cudaHostRegister; rank 0, informational
    e2e_pageable = None
    if rank == 0 and match:
        pg_in = np.array(hnp, copy=True)
        pg = (np.zeros_like(kp_out), np.zeros_like(de_out), np.zeros_like(cnt_out), np.zeros_like(m12_out), np.zeros_like(nm_out))
        ex.extract_match_to_host(pg_in, *pg, WINDOW, NNRATIO, True)
        t0 = time.perf_counter()
        for _ in range(max(1, args.steps // 2)):
            ex.extract_match_to_host(pg_in, *pg, WINDOW, NNRATIO, True)
        torch.cuda.synchronize()
        e2e_pageable = F * max(1, args.steps // 2) / (time.perf_counter() - t0)
        del pg_in, pg
    h2d = F * W * H
    # the same pinned input copied alone: the PCIe bound of the end-to-end number (images are the input, they must cross)
    pcie_dst = torch.empty((F, H, W), dtype=torch.uint8, device=dev)
    pcie_dst.copy_(host, non_blocking=True)
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    pcie_dst.copy_(host, non_blocking=True)
    ev1.record()
    torch.cuda.synchronize()
    h2d_gbs = h2d / (ev0.elapsed_time(ev1) * 1e-3) / 1e9
    del pcie_dst
    d2h = kp_out.nbytes + de_out.nbytes + cnt_out.nbytes + m12_out.nbytes + nm_out.nbytes + 8
    # descriptor pairs the guided search of one step evaluates (the kernels count them on request; outside the timed region)
    if match:
        ncmp.value = ex.search_init_pairs(0, F - 1, WINDOW, NNRATIO, True, fetch=True)[2]

    # ---- roofline of the dominant kernel ----------------------------------------------------------
    peak, peak_src = measured_peaks()
    ab = algorithmic_bytes(W, H, nkp)
    frames_total = F * args.steps
    ext_stages = {k: v for k, v in stage_ms.items() if v > 0}
    dom = max(ext_stages, key=ext_stages.get) if ext_stages else "fast"
    dom_bytes = {"resize": ab["resize"], "blur": ab["blur"], "fast": ab["fast"], "describe": ab["describe"]}.get(dom, ab["fast"])
    dom_launches = max(1, stage_launches.get(dom, 1))
    dom_ms = stage_ms.get(dom, 0.0)
    achieved = (dom_bytes * frames_total / dom_launches) / (dom_ms / dom_launches * 1e-3) / 1e9 if dom_ms > 0 else 0.0
    # DRAM traffic and executed warp instructions of the dominant kernel: dram__bytes_read.sum + dram__bytes_write.sum
    # and smsp__inst_executed.sum of one `ncu --set full` capture of the 512-frame chunk this run launches, per frame
    # (profiles/ncu_constants.json), scaled to the frames of one launch.  null when the capture is stale or absent.
    consts, why = ncu_constants()
    traffic = None
    if consts and (W, H) == (640, 480) and dom in consts["kernels"]:
        traffic = consts["kernels"][dom]["dram_bytes_per_frame"] * frames_total / dom_launches
    roofline = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic,
                "traffic_source": (f"ncu --set full of a {consts['chunk_frames']}-frame launch, {consts['source']}" if traffic else why),
                "algorithmic_bytes_per_launch": dom_bytes * frames_total / dom_launches,
                "peak_source": peak_src,
                "algorithmic_bytes_per_frame": dom_bytes,
                "share_of_device_time": dom_ms / max(sum(stage_ms.values()), 1e-9)}
    if consts and (W, H) == (640, 480) and dom in consts["kernels"] and clk.get("sm_mhz") and dom_ms > 0:
        # the bound that actually limits this kernel: issue slots (148 SMs x 4 schedulers x SM clock under load)
        wi = consts["kernels"][dom]["warp_inst_per_frame"]
        ginst = wi * (frames_total / dom_launches) / (dom_ms / dom_launches * 1e-3) / 1e9
        issue_peak = 148 * 4 * clk["sm_mhz"] * 1e6 / 1e9
        roofline["issue"] = {"achieved": ginst, "peak": issue_peak, "unit": "G warp-inst/s", "frac": ginst / issue_peak,
                             "warp_inst_per_frame": wi, "source": f"smsp__inst_executed.sum, {consts['source']}"}
    path_gbs = ab["total"] * value / world / 1e9
    stages = {k: {"ms_per_step": v / args.steps, "launches_per_step": stage_launches[k] / args.steps}
              for k, v in stage_ms.items() if stage_launches[k]}

    line = {
        "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": workload_name,
                   "frames_per_step_per_gpu": F, "chunk_frames": args.chunk, "e2e_chunk_frames": min(args.chunk, int(os.environ.get("ORBX_HOST_CHUNK", "256"))), "keypoints_per_frame": nkp,
                   "l2_policy": f"inputs ({F * W * H / 1e9:.2f} GB per step) larger than L2", "sharding": "frames by rank, no collective", "rank_cpu_binding": numa},
        "clocks": clk,
        "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "h2d_gbs_measured": h2d_gbs, "pcie_bound_frames_s": world * h2d_gbs * 1e9 / (W * H),
                "pageable_buffers_frames_s_rank0": e2e_pageable},
        "gpu_launches": int(sum(stage_launches.values())),
        "roofline": roofline,
        "path_roofline": {"algorithmic_bytes_per_frame": ab["total"], "achieved": path_gbs, "peak": peak,
                          "frac": path_gbs / peak, "unit": "GB/s"},
        "stages": stages,
        "matches_per_pair": float(nm_out.mean()), "hamming_compares_per_step": int(ncmp.value),
    }

    # ---- second half of the metric: Hamming cmp/s of the brute-force 2-NN kernel on this GPU (short run) ----
    # measured on every rank of this run through orbm_knn2_sharded (NCCL at N > 1), config 5 as named
    if args.workload == "frames" and not args.no_hamming:
        line["hamming"] = hamming_leg(dev, rank, world, stream, cpu=not args.no_cpu)

    # ---- CPU baseline (rank 0, N=1 only): bounded sample of the same frames ----------------------
    if rank == 0 and world == 1 and not args.no_cpu:
        cores = os.cpu_count() or 1
        s1 = 16
        sN = min(max(16 * cores, 64), F)      # 16 frames per thread: two frames per thread measured 40 % low (thread start-up and imbalance)
        win = WINDOW if match else 0          # window 0: the matcher finds no candidate (extraction-only workloads)
        if args.workload == "hd":
            s1, sN = 4, max(cores, 8)
        tN, kind = cpu_extract_match(hnp[:sN], cores, True, win)
        t1, _ = cpu_extract_match(hnp[:s1], 1, True, win)
        cb = {"value": sN / tN, "unit": "frames/s", "cores": cores, "kind": kind,
              "sample": f"first {sN} frames (+{sN - 1} pairs), frames dealt over {cores} threads; "
                        f"single-thread on {s1} frames: {s1 / t1:.2f} frames/s",
              "single_thread_value": s1 / t1}
        # "as the reference compiles it": the same unmodified ORBextractor.cc at -O3 -march=<this host's level> (timing only)
        tF, _ = cpu_extract_match(hnp[:sN], cores, True, win, fast=True)
        if tF:
            cb["value_o3_native"] = sN / tF
            cb["o3_native_build"] = f"-O3 -march=x86-64-{cpu_level()} (oracle/Makefile ref_fast), {cores} threads, same sample"
        if args.workload == "frames":
            # single thread, one operator() call per frame, median over 200 frames
            n1 = min(200, F)
            for key, fast in (("single_thread_median_ms", False), ("single_thread_median_ms_o3_native", True)):
                ts = cpu_single_thread_ms(hnp[:n1], fast)
                if ts:
                    cb[key] = float(np.median(ts)); cb[key + "_frames"] = n1
            prim = cv2_primitives_ms(hnp[:min(100, F)])
            if prim:
                cb["cv2_primitives_ms"] = prim
        line["cpu_baseline"] = cb
    if rank == 0:
        emit(line)
    ex.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
