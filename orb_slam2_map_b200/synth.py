"""Deterministic synthetic inputs for tests and bench (SURVEY.md §8(d) generator).

Everything is integer / exactly-rounded arithmetic keyed by a counter hash, so the same bytes come
out on any machine (numpy here, torch on the GPU for the big batches).  The reference ships no
images (TUM/KITTI data are not in the repo), so these stand in for camera frames:

* ``textured_canvas``  — 4 octaves of bilinear value noise + random rectangles, 8-bit.
* ``frame_sequence``   — frame k = window of the canvas shifted by (k, k//2) px plus +-2 noise, the
  "slowly translating camera" that ``ORBmatcher::SearchForInitialization`` (window 100) can match.
* ``stereo_pair``      — right image = left shifted by a piecewise-constant disparity field.
* ``random_descriptors`` — iid 256-bit descriptors (+ planted near-duplicates) for the 2-NN search.
"""
from __future__ import annotations

import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(x: np.ndarray) -> np.ndarray:
    """splitmix64 finaliser on uint64 arrays (wrap-around arithmetic)."""
    with np.errstate(over="ignore"):
        z = (x.astype(np.uint64) + np.uint64(0x9E3779B97F4A7C15)) & _M64
        z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
        z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
        return z ^ (z >> np.uint64(31))


def _u01(seed: int, stream: int, n: int) -> np.ndarray:
    """n uniform doubles in [0,1) from (seed, stream)."""
    ctr = np.arange(n, dtype=np.uint64) + (np.uint64(seed) << np.uint64(40)) + (np.uint64(stream) << np.uint64(32))
    return (splitmix64(ctr) >> np.uint64(11)).astype(np.float64) / float(1 << 53)


def textured_canvas(width: int, height: int, seed: int = 0, nrect: int | None = None) -> np.ndarray:
    """8-bit textured image: value noise (lattices 8,16,32,64 px) + rectangles, min-max normalised."""
    acc = np.zeros((height, width), dtype=np.float64)
    for o, cell in enumerate((8, 16, 32, 64)):
        gw, gh = width // cell + 2, height // cell + 2
        lat = _u01(seed, 1 + o, gw * gh).reshape(gh, gw) * 255.0
        ys, xs = np.arange(height), np.arange(width)
        y0, x0 = ys // cell, xs // cell
        fy = ((ys % cell).astype(np.float64) / cell)[:, None]
        fx = ((xs % cell).astype(np.float64) / cell)[None, :]
        a = lat[y0][:, x0]
        b = lat[y0][:, x0 + 1]
        c = lat[y0 + 1][:, x0]
        d = lat[y0 + 1][:, x0 + 1]
        top = a + (b - a) * fx
        bot = c + (d - c) * fx
        acc += (top + (bot - top) * fy) * (cell / 64.0 + 0.25)
    if nrect is None:
        nrect = max(1, int(300 * (width * height) / (640.0 * 480.0)))
    r = _u01(seed, 9, nrect * 16).reshape(nrect, 16)
    for i in range(nrect):
        w = 4 + int(r[i, 0] * 37)
        h = 4 + int(r[i, 1] * 37)
        x = int(r[i, 2] * max(1, width - w))
        y = int(r[i, 3] * max(1, height - h))
        amp = (r[i, 4:16].sum() - 6.0) * 40.0      # Irwin-Hall ~ N(0,40)
        acc[y:y + h, x:x + w] += amp
    lo, hi = acc.min(), acc.max()
    out = np.rint((acc - lo) * (255.0 / (hi - lo)))
    return out.astype(np.uint8)


def _noise_pm2(seed: int, k: int, n: int) -> np.ndarray:
    ctr = np.arange(n, dtype=np.uint64) + (np.uint64(seed + 1) << np.uint64(44)) + (np.uint64(k) << np.uint64(24))
    return (splitmix64(ctr) % np.uint64(5)).astype(np.int16) - 2


def frame_from_canvas(canvas: np.ndarray, k: int, width: int, height: int, seed: int = 0) -> np.ndarray:
    """Frame k of the translating sequence (offset wraps inside the canvas)."""
    H, W = canvas.shape
    ox = k % (W - width + 1)
    oy = (k // 2) % (H - height + 1)
    win = canvas[oy:oy + height, ox:ox + width].astype(np.int16)
    win = win + _noise_pm2(seed, k, width * height).reshape(height, width)
    return np.clip(win, 0, 255).astype(np.uint8)


def frame_sequence(nframes: int, width: int, height: int, seed: int = 0, span: int = 256) -> np.ndarray:
    """[nframes, height, width] uint8; canvas is (width+span) x (height+span//2+1)."""
    canvas = textured_canvas(width + span, height + span // 2 + 1, seed)
    return np.stack([frame_from_canvas(canvas, k, width, height, seed) for k in range(nframes)])


def stereo_pair(width: int, height: int, seed: int = 0, k: int = 0):
    """(left, right): right(x) = left(x + d(y-band)), disparities 5..60 px in 8 horizontal bands."""
    canvas = textured_canvas(width + 128, height, seed)
    left = canvas[:, 64:64 + width].copy()
    right = np.empty_like(left)
    nb = 8
    disp = 5 + (splitmix64(np.arange(nb, dtype=np.uint64) + np.uint64(1000 * seed + 7)) % np.uint64(56)).astype(np.int64)
    for b in range(nb):
        y0, y1 = b * height // nb, (b + 1) * height // nb
        d = int(disp[b])
        right[y0:y1] = canvas[y0:y1, 64 + d:64 + d + width]
    n = _noise_pm2(seed + 17, k, width * height).reshape(height, width)
    right = np.clip(right.astype(np.int16) + n, 0, 255).astype(np.uint8)
    return left, right


def random_descriptors(n: int, seed: int) -> np.ndarray:
    """[n, 32] uint8 iid uniform bits."""
    ctr = np.arange(n * 4, dtype=np.uint64) + (np.uint64(seed) << np.uint64(48))
    return splitmix64(ctr).view(np.uint8).reshape(n, 32).copy()


def plant_near_duplicates(queries: np.ndarray, database: np.ndarray, frac: float = 0.01, seed: int = 3,
                          max_flips: int = 20) -> np.ndarray:
    """Overwrite a fraction of queries with database rows having <= max_flips flipped bits."""
    q = queries.copy()
    nq, nd = len(q), len(database)
    m = max(1, int(nq * frac))
    h = splitmix64(np.arange(m * 2, dtype=np.uint64) + (np.uint64(seed) << np.uint64(50)))
    qi = (h[:m] % np.uint64(nq)).astype(np.int64)
    di = (h[m:] % np.uint64(nd)).astype(np.int64)
    q[qi] = database[di]
    fl = splitmix64(np.arange(m * max_flips, dtype=np.uint64) + (np.uint64(seed) << np.uint64(52))).reshape(m, max_flips)
    for j in range(max_flips):
        bit = (fl[:, j] % np.uint64(256)).astype(np.int64)
        use = ((fl[:, j] >> np.uint64(20)) & np.uint64(1)).astype(bool)
        rows = qi[use]
        b = bit[use]
        q[rows, b // 8] ^= (1 << (b % 8)).astype(np.uint8)
    return q


# ---- torch twin of frame_sequence (same bytes), used by bench.py to build big batches on the GPU ----
def _lsr64(z, s):
    return (z >> s) & ((1 << (64 - s)) - 1)


def _splitmix64_torch(x):
    """splitmix64 on int64 tensors holding uint64 bit patterns (two's-complement wrap-around)."""
    z = x + (-7046029254386353131)                    # 0x9E3779B97F4A7C15
    z = (z ^ _lsr64(z, 30)) * (-4658895280553007687)   # 0xBF58476D1CE4E5B9
    z = (z ^ _lsr64(z, 27)) * (-7723592293110705685)   # 0x94D049BB133111EB
    return z ^ _lsr64(z, 31)


def frame_sequence_torch(nframes: int, width: int, height: int, seed: int = 0, span: int = 256,
                         device="cuda", out=None, first: int = 0):
    """[nframes, height, width] uint8 torch tensor, byte-identical to frame_sequence()[first:first+nframes]."""
    import torch
    canvas_np = textured_canvas(width + span, height + span // 2 + 1, seed)
    canvas = torch.from_numpy(canvas_np).to(device)
    H, W = canvas_np.shape
    if out is None:
        out = torch.empty((nframes, height, width), dtype=torch.uint8, device=device)
    idx = torch.arange(width * height, dtype=torch.int64, device=device)
    for i in range(nframes):
        k = first + i
        ox = k % (W - width + 1)
        oy = (k // 2) % (H - height + 1)
        z = _splitmix64_torch(idx + (((seed + 1) << 44) + (k << 24)))
        m = torch.remainder(torch.remainder(z, 5) + (z < 0).to(torch.int64), 5).to(torch.int16) - 2
        win = canvas[oy:oy + height, ox:ox + width].to(torch.int16) + m.view(height, width)
        out[i] = win.clamp_(0, 255).to(torch.uint8)
    return out
