"""CPU tests (no GPU) of the matcher half of the oracle: each routine against an independent
numpy / pure-Python statement of the reference semantics on small inputs, plus domain properties.
(The reference ships no tests or vectors for src/ORBmatcher.cc / src/Frame.cc — SURVEY.md §4.)"""
import math

import numpy as np

import oracle_lib as O
from orb_slam2_map_b200 import synth


def _popcount_rows(a, b):
    return np.unpackbits(a ^ b, axis=-1).sum(axis=-1)


def test_descriptor_distance_kats():
    rng = np.random.default_rng(0)
    a = rng.integers(0, 256, (512, 32), dtype=np.uint8)
    b = rng.integers(0, 256, (512, 32), dtype=np.uint8)
    exp = _popcount_rows(a, b)
    for i in range(len(a)):
        assert O.descriptor_distance(a[i], b[i]) == exp[i]
    z, f = np.zeros(32, np.uint8), np.full(32, 255, np.uint8)
    assert O.descriptor_distance(z, z) == 0 and O.descriptor_distance(z, f) == 256
    one = z.copy(); one[17] = 0x10
    assert O.descriptor_distance(z, one) == 1


def _rand_frame(n, seed, w=640, h=480, nlev=3, bounds=None):
    r = np.random.default_rng(seed)
    k = np.zeros(n, O.KP_DTYPE)
    k["x"] = r.uniform(0, w, n).astype(np.float32)
    k["y"] = r.uniform(0, h, n).astype(np.float32)
    k["octave"] = r.integers(0, nlev, n)
    k["angle"] = r.uniform(0, 360, n).astype(np.float32)
    k["size"] = 31
    k["class_id"] = -1
    d = r.integers(0, 256, (n, 32), dtype=np.uint8)
    return O.FrameData(k, d, w, h, bounds=bounds)


def _features_in_area_py(f, x, y, r, min_level, max_level):
    """Frame::GetFeaturesInArea (reference src/Frame.cc:327-380) + AssignFeaturesToGrid (:230-245)
    + PosInGrid (:382-392), written independently in float32 numpy."""
    f32 = np.float32
    minx, maxx, miny, maxy = (f32(v) for v in f.bounds)
    inv_w, inv_h = f32(64) / (maxx - minx), f32(48) / (maxy - miny)
    grid = [[[] for _ in range(48)] for _ in range(64)]
    for i, kp in enumerate(f.kps):
        px = int(np.round(f32(f32(kp["x"]) - minx) * inv_w))      # C round(): half away from zero
        py = int(np.round(f32(f32(kp["y"]) - miny) * inv_h))
        fx = float(f32(f32(kp["x"]) - minx) * inv_w); fy = float(f32(f32(kp["y"]) - miny) * inv_h)
        px = int(math.floor(fx + 0.5)) if fx >= 0 else -int(math.floor(-fx + 0.5))
        py = int(math.floor(fy + 0.5)) if fy >= 0 else -int(math.floor(-fy + 0.5))
        if 0 <= px < 64 and 0 <= py < 48:
            grid[px][py].append(i)
    x, y, r = f32(x), f32(y), f32(r)
    c0 = max(0, int(math.floor(float(f32(f32(x - minx) - r) * inv_w))))
    if c0 >= 64:
        return []
    c1 = min(63, int(math.ceil(float(f32(f32(x - minx) + r) * inv_w))))
    if c1 < 0:
        return []
    r0 = max(0, int(math.floor(float(f32(f32(y - miny) - r) * inv_h))))
    if r0 >= 48:
        return []
    r1 = min(47, int(math.ceil(float(f32(f32(y - miny) + r) * inv_h))))
    if r1 < 0:
        return []
    check = (min_level > 0) or (max_level >= 0)
    out = []
    for ix in range(c0, c1 + 1):
        for iy in range(r0, r1 + 1):
            for i in grid[ix][iy]:
                kp = f.kps[i]
                if check:
                    if kp["octave"] < min_level:
                        continue
                    if max_level >= 0 and kp["octave"] > max_level:
                        continue
                if abs(f32(kp["x"]) - x) < r and abs(f32(kp["y"]) - y) < r:
                    out.append(i)
    return out


def test_features_in_area_vs_python():
    for seed, bounds in [(1, None), (2, (-12.5, 655.0, -8.25, 491.0))]:
        f = _rand_frame(700, seed, bounds=bounds)
        r = np.random.default_rng(seed + 10)
        for _ in range(150):
            x, y = r.uniform(-30, 670), r.uniform(-30, 510)
            rad = float(r.choice([3.0, 10.0, 47.5, 100.0, 400.0]))
            lo, hi = [(-1, -1), (0, 0), (1, 2), (0, -1), (2, -1), (1, 0)][int(r.integers(0, 6))]
            got = O.features_in_area(f, x, y, rad, lo, hi).tolist()
            assert got == _features_in_area_py(f, x, y, rad, lo, hi), (x, y, rad, lo, hi)


def _three_maxima(h):
    m1 = m2 = m3 = 0
    i1 = i2 = i3 = -1
    for i, s in enumerate(h):
        if s > m1:
            m3, m2, m1 = m2, m1, s; i3, i2, i1 = i2, i1, i
        elif s > m2:
            m3, m2 = m2, s; i3, i2 = i2, i
        elif s > m3:
            m3 = s; i3 = i
    if m2 < np.float32(0.1) * np.float32(m1):
        i2 = i3 = -1
    elif m3 < np.float32(0.1) * np.float32(m1):
        i3 = -1
    return i1, i2, i3


def _search_init_py(f1, f2, prev, window, nnratio, check_ori):
    """ORBmatcher::SearchForInitialization (reference src/ORBmatcher.cc:405-520), pure Python."""
    n1, n2 = len(f1.kps), len(f2.kps)
    m12 = [-1] * n1
    m21 = [-1] * n2
    md = [2 ** 31 - 1] * n2
    hist = [[] for _ in range(30)]
    nm = 0
    f32 = np.float32
    for i1 in range(n1):
        if f1.kps[i1]["octave"] > 0:
            continue
        cand = _features_in_area_py(f2, prev[i1][0], prev[i1][1], window, 0, 0)
        if not cand:
            continue
        best, best2, bi = 2 ** 31 - 1, 2 ** 31 - 1, -1
        for i2 in cand:
            d = int(_popcount_rows(f1.desc[i1], f2.desc[i2]))
            if md[i2] <= d:
                continue
            if d < best:
                best2, best, bi = best, d, i2
            elif d < best2:
                best2 = d
        if best <= 50 and f32(best) < f32(best2) * f32(nnratio):
            if m21[bi] >= 0:
                m12[m21[bi]] = -1
                nm -= 1
            m12[i1] = bi; m21[bi] = i1; md[bi] = best; nm += 1
            if check_ori:
                rot = f32(f1.kps[i1]["angle"]) - f32(f2.kps[bi]["angle"])
                if rot < 0:
                    rot = f32(rot + f32(360.0))
                v = float(f32(rot * f32(1.0 / 30)))
                b = int(math.floor(v + 0.5))
                if b == 30:
                    b = 0
                hist[b].append(i1)
    if check_ori:
        keep = _three_maxima([len(h) for h in hist])
        for b in range(30):
            if b in keep:
                continue
            for i1 in hist[b]:
                if m12[i1] >= 0:
                    m12[i1] = -1; nm -= 1
    return nm, m12


def test_search_for_initialization_vs_python():
    """Few distinct descriptors -> many ties, steals and vMatchedDistance skips."""
    for seed in (3, 4):
        f1 = _rand_frame(260, seed, nlev=2)
        f2 = _rand_frame(300, seed + 50, nlev=2)
        base = np.random.default_rng(seed).integers(0, 256, (6, 32), dtype=np.uint8)
        r = np.random.default_rng(seed + 1)
        for f in (f1, f2):
            d = base[r.integers(0, 6, len(f.kps))].copy()
            d[np.arange(len(d)), r.integers(0, 32, len(d))] ^= (1 << r.integers(0, 8, len(d))).astype(np.uint8)
            f.desc[:] = d
        prev = np.stack([f1.kps["x"], f1.kps["y"]], axis=1).astype(np.float32)
        for window, ratio, ori in [(100, 0.9, True), (60, 0.99, False), (200, 0.99, True)]:
            n_o, m_o, p_o, ncmp = O.search_for_initialization(f1, f2, prev, window, ratio, ori)
            n_p, m_p = _search_init_py(f1, f2, prev, window, ratio, ori)
            assert n_o == n_p and m_o.tolist() == m_p, (seed, window)
            assert n_o == int((m_o >= 0).sum()) and ncmp > 0
            # matches are one-to-one and prev_matched is updated only for matched keypoints
            used = m_o[m_o >= 0]
            assert len(set(used.tolist())) == len(used)
            un = m_o < 0
            assert np.array_equal(p_o[un], prev[un])
            assert np.array_equal(p_o[~un, 0], f2.kps["x"][m_o[~un]])


def test_search_init_on_extracted_sequence_finds_true_motion():
    frames = synth.frame_sequence(2, 640, 480, seed=5)
    ex = O.Extractor()
    (k1, d1), (k2, d2) = ex.extract(frames[0]), ex.extract(frames[1])
    f1, f2 = O.FrameData(k1, d1, 640, 480), O.FrameData(k2, d2, 640, 480)
    prev = np.stack([k1["x"], k1["y"]], axis=1)
    n, m12, _, _ = O.search_for_initialization(f1, f2, prev, 100, 0.9, True)
    assert n > 40
    sel = m12 >= 0
    dx = k2["x"][m12[sel]] - k1["x"][sel]
    dy = k2["y"][m12[sel]] - k1["y"][sel]
    # frame 1 = canvas window shifted by (1, 0) px => features move by (-1, 0)
    assert abs(float(np.median(dx)) + 1.0) < 0.6 and abs(float(np.median(dy))) < 0.6


def test_knn2_vs_numpy_and_ties():
    q = synth.random_descriptors(97, 1)
    db = synth.random_descriptors(1500, 2)
    db[700] = db[20]                      # exact duplicates: lower index must win
    q[5] = db[20]
    q = synth.plant_near_duplicates(q, db, frac=0.1)
    q[5] = db[20]
    od, oi = O.knn2(q, db, offset=1000, nthreads=3)
    D = _popcount_rows(q[:, None, :], db[None, :, :])
    for i in range(len(q)):
        order = np.lexsort((np.arange(len(db)), D[i]))
        assert oi[i, 0] == 1000 + order[0] and oi[i, 1] == 1000 + order[1]
        assert od[i, 0] == D[i, order[0]] and od[i, 1] == D[i, order[1]]
    assert oi[5, 0] == 1020 and oi[5, 1] == 1700 and od[5, 0] == 0 and od[5, 1] == 0
    # sharded == whole: merging per-shard top-2 with (dist, index) order gives the same answer
    parts = [O.knn2(q, db[s:s + 500], offset=1000 + s) for s in (0, 500, 1000)]
    for i in range(len(q)):
        c = sorted((int(pd[i, j]), int(pi[i, j])) for pd, pi in parts for j in range(2))
        assert (c[0][1], c[1][1]) == (oi[i, 0], oi[i, 1])


def test_distribute_octtree_properties():
    rng = np.random.default_rng(7)
    for n, N, (w, h) in [(3000, 217, (608, 448)), (500, 300, (608, 448)), (40, 100, (300, 200)), (4000, 434, (1209, 344)), (1, 5, (100, 100))]:
        xyr = np.stack([rng.integers(0, w, n), rng.integers(0, h, n), rng.integers(7, 120, n)], axis=1).astype(np.float32)
        xyr = np.unique(xyr[:, :2], axis=0, return_index=True)[1]
        pts = np.stack([rng.integers(0, w, n), rng.integers(0, h, n), rng.integers(7, 120, n)], axis=1).astype(np.float32)
        idx = O.distribute_octtree(pts, 16, 16 + w, 16, 16 + h, N)
        assert len(set(idx.tolist())) == len(idx)                     # no key twice
        assert len(idx) <= max(N + 3, 4 * max(1, round(w / h)))       # SURVEY.md App. C.4 bound
        if n >= 4 * N:
            assert len(idx) >= N
        if n <= 1:
            assert len(idx) == n
    # all keys identical -> they can never be separated: one survivor, the max response (first wins)
    same = np.tile(np.array([[50, 60, 10]], np.float32), (30, 1))
    same[7, 2] = 99; same[9, 2] = 99
    assert O.distribute_octtree(same, 16, 624, 16, 464, 100).tolist() == [7]


def test_compute_stereo_matches_recovers_disparity():
    w, h = 640, 240
    left, right = synth.stereo_pair(w, h, seed=2)
    el, er = O.Extractor(800), O.Extractor(800)
    kl, dl = el.extract(left)
    kr, dr = er.extract(right)
    t = el.tables()
    pl = [el.level(l) for l in range(8)]
    pr = [er.level(l) for l in range(8)]
    bf, fx = 386.1448, 718.856
    ur, depth, ncmp = O.compute_stereo_matches(kl, dl, kr, dr, pl, pr, t["scale"], t["inv_scale"], bf, fx)
    ok = ur >= 0
    assert ok.sum() > 100 and ncmp > ok.sum()
    assert np.all(depth[~ok] == -1) and np.all(depth[ok] > 0)
    disp = kl["x"][ok] - ur[ok]
    assert np.allclose(depth[ok], np.float32(bf) / disp, rtol=1e-6)
    # synthetic disparity field: 8 horizontal bands of constant disparity in 5..60
    band = (kl["y"][ok] * 8 / h).astype(int)
    for b in range(8):
        d = disp[band == b]
        if len(d) >= 8:
            med = float(np.median(d))
            assert 4.0 <= med <= 61.0
            assert np.mean(np.abs(d - med) < 1.5) > 0.5      # band edges and coarse octaves straddle bands
