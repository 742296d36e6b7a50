"""oracle/kfmatch_oracle.cpp (SearchForTriangulation, Fuse x2, SearchBySim3) against independent Python statements
of reference src/ORBmatcher.cc:657-823, 825-973, 975-1100, 1102-1326 in float32 numpy."""
import numpy as np

import oracle_lib as O
from test_oracle_matchers import _features_in_area_py, _three_maxima

f32 = np.float32
NLEV = 8
SCALE = (f32(1.2) ** np.arange(NLEV)).astype(np.float32)
SIGMA2 = (SCALE * SCALE).astype(np.float32)


def _dist(a, b):
    return int(np.unpackbits(np.bitwise_xor(a, b)).sum())


def flip_bits(rng, d, nbits):
    d = d.copy()
    for b in rng.choice(256, nbits, replace=False):
        d[b >> 3] ^= np.uint8(1 << (b & 7))
    return d


def make_tri_pair(seed, n1=400, n2=520, nodes=24, mono_frac=0.5):
    """KF1 random; KF2 = resampled KF1 features moved along x with small y jitter (so horizontal epipolar lines
    accept some and reject others), duplicated descriptors for distance ties, plus distractors."""
    rng = np.random.default_rng(seed)
    k1 = np.zeros(n1, O.KP_DTYPE)
    k1["x"] = rng.uniform(20, 620, n1).astype(np.float32)
    k1["y"] = rng.uniform(20, 460, n1).astype(np.float32)
    k1["octave"] = rng.integers(0, NLEV, n1)
    k1["angle"] = rng.uniform(0, 360, n1).astype(np.float32)
    d1 = rng.integers(0, 256, (n1, 32), dtype=np.uint8)
    node1 = rng.integers(0, nodes, n1).astype(np.uint32)
    src = rng.integers(0, n1, n2)
    k2 = k1[src].copy()
    k2["x"] = (k2["x"] + rng.uniform(-40, 40, n2)).astype(np.float32)
    k2["y"] = (k2["y"] + rng.normal(0, 1.2, n2)).astype(np.float32)
    k2["octave"] = np.clip(k2["octave"] + rng.integers(-1, 2, n2), 0, NLEV - 1)
    k2["angle"] = ((k2["angle"] + np.where(rng.random(n2) < 0.8, rng.normal(12, 3, n2), rng.uniform(0, 360, n2))) % 360).astype(np.float32)
    d2 = np.stack([flip_bits(rng, d1[s], int(rng.choice([0, 0, 3, 10, 30, 60]))) for s in src])
    node2 = node1[src].copy()
    # a cluster around the epipole, so the closeness test (:744-750) rejects some mono-mono candidates
    ex, ey = f32(322.5), f32(241.25)
    near = rng.choice(n2, n2 // 6, replace=False)
    k2["x"][near] = (ex + rng.uniform(-14, 14, len(near))).astype(np.float32)
    k2["y"][near] = (k1["y"][src[near]] + rng.normal(0, 0.5, len(near))).astype(np.float32)
    ur1 = np.where(rng.random(n1) < mono_frac, -1, k1["x"] - 5).astype(np.float32)
    ur2 = np.where(rng.random(n2) < mono_frac, -1, k2["x"] - 5).astype(np.float32)
    fl1 = (1 * (rng.random(n1) < 0.9) + 2 * (rng.random(n1) < 0.15)).astype(np.uint8)
    fl2 = (1 * (rng.random(n2) < 0.9) + 2 * (rng.random(n2) < 0.15)).astype(np.uint8)
    kf1 = O.KeyFrameData(k1, d1, 640, 480, ur1, SCALE, SIGMA2, node1, fl1)
    kf2 = O.KeyFrameData(k2, d2, 640, 480, ur2, SCALE, SIGMA2, node2, fl2)
    # lines l = x1' F12 ~ (0, 1, -y1) plus a small general part so every term of :143-145 matters
    F = np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], np.float32) + rng.normal(0, 2e-4, (3, 3)).astype(np.float32)
    return kf1, kf2, F.reshape(9), (ex, ey)


def tri_py(kf1, kf2, F, epi, only_stereo, check_ori):
    A, B = kf1.fd, kf2.fd
    F = np.asarray(F, np.float32).reshape(3, 3)
    ex, ey = f32(epi[0]), f32(epi[1])
    fv1, fv2 = {}, {}
    for i in range(A.c.n):
        if kf1.flags[i] & 1:
            fv1.setdefault(int(kf1.node[i]), []).append(i)
    for i in range(B.c.n):
        if kf2.flags[i] & 1:
            fv2.setdefault(int(kf2.node[i]), []).append(i)
    m12 = np.full(A.c.n, -1, np.int32)
    hist = [[] for _ in range(30)]
    n = 0
    for node in sorted(set(fv1) & set(fv2)):
        for i1 in fv1[node]:
            if kf1.flags[i1] & 2:
                continue
            st1 = A.u_right[i1] >= 0
            if only_stereo and not st1:
                continue
            x1, y1 = f32(A.kps["x"][i1]), f32(A.kps["y"][i1])
            a = f32(f32(x1 * F[0, 0]) + f32(y1 * F[1, 0])) + F[2, 0]
            b = f32(f32(x1 * F[0, 1]) + f32(y1 * F[1, 1])) + F[2, 1]
            c = f32(f32(x1 * F[0, 2]) + f32(y1 * F[1, 2])) + F[2, 2]
            best, bi = 50, -1
            for i2 in fv2[node]:
                if kf2.flags[i2] & 2:
                    continue
                st2 = B.u_right[i2] >= 0
                if only_stereo and not st2:
                    continue
                d = _dist(A.desc[i1], B.desc[i2])
                if d > 50 or d > best:
                    continue
                x2, y2, o2 = f32(B.kps["x"][i2]), f32(B.kps["y"][i2]), int(B.kps["octave"][i2])
                if not st1 and not st2:
                    dx, dy = f32(ex - x2), f32(ey - y2)
                    if f32(f32(dx * dx) + f32(dy * dy)) < f32(f32(100) * SCALE[o2]):
                        continue
                num = f32(f32(f32(a * x2) + f32(b * y2)) + c)
                den = f32(f32(a * a) + f32(b * b))
                if den == 0:
                    continue
                dsqr = f32(f32(num * num) / den)
                if float(dsqr) < 3.84 * float(SIGMA2[o2]):
                    bi, best = i2, d
            if bi >= 0:
                m12[i1] = bi
                n += 1
                if check_ori:
                    rot = f32(f32(A.kps["angle"][i1]) - f32(B.kps["angle"][bi]))
                    if rot < 0:
                        rot = f32(rot + f32(360))
                    v = float(f32(rot * f32(f32(1.0) / f32(30))))
                    bn = int(np.floor(v + 0.5))
                    if bn == 30:
                        bn = 0
                    hist[bn].append(i1)
    if check_ori:
        i1_, i2_, i3_ = _three_maxima([len(h) for h in hist])
        for bn in range(30):
            if bn in (i1_, i2_, i3_):
                continue
            for j in hist[bn]:
                m12[j] = -1
                n -= 1
    return n, m12


def test_search_for_triangulation_vs_python():
    total = 0
    for seed, only_stereo, ori, mono in [(1, False, True, 0.5), (2, True, True, 0.3), (3, False, False, 1.0), (4, False, True, 0.0)]:
        kf1, kf2, F, epi = make_tri_pair(seed, mono_frac=mono)
        n_o, m_o = O.search_for_triangulation(kf1, kf2, F, epi, only_stereo, ori)
        n_p, m_p = tri_py(kf1, kf2, F, epi, only_stereo, ori)
        assert n_o == n_p and np.array_equal(m_o, m_p)
        assert n_o == int((m_o >= 0).sum())
        total += n_o
    assert total > 30          # the scenes do produce matches


def make_fuse_case(seed, n=900, nq=500, th=3.0):
    rng = np.random.default_rng(seed)
    k = np.zeros(n, O.KP_DTYPE)
    k["x"] = rng.uniform(0, 640, n).astype(np.float32)
    k["y"] = rng.uniform(0, 480, n).astype(np.float32)
    # clumps: many keypoints share a search window, with duplicated descriptors for ties across grid cells
    cl = rng.choice(n, n // 3, replace=False)
    k["x"][cl] = (k["x"][cl[0]] + rng.uniform(-25, 25, len(cl))).astype(np.float32) % 640
    k["y"][cl] = (k["y"][cl[0]] + rng.uniform(-25, 25, len(cl))).astype(np.float32) % 480
    k["octave"] = rng.integers(0, NLEV, n)
    d = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    d[cl[1::2]] = d[cl[0]]
    ur = np.where(rng.random(n) < 0.5, -1, k["x"] - 4).astype(np.float32)
    kf = O.KeyFrameData(k, d, 640, 480, ur, SCALE, SIGMA2)
    src = rng.integers(0, n, nq)
    q = np.zeros(nq, O.PROJ_DTYPE)
    q["u"] = (k["x"][src] + rng.normal(0, 1.5, nq)).astype(np.float32)
    q["v"] = (k["y"][src] + rng.normal(0, 1.5, nq)).astype(np.float32)
    lvl = np.clip(k["octave"][src] + rng.integers(0, 2, nq), 0, NLEV - 1)
    q["radius"] = (f32(th) * SCALE[lvl]).astype(np.float32)
    q["u_right"] = (q["u"] - 4 + rng.normal(0, 1.0, nq)).astype(np.float32)
    q["level_lo"] = lvl - 1
    q["level_hi"] = lvl
    q["min_level"] = -1
    q["max_level"] = -1
    # a few queries off the image / far from everything
    q["u"][:5] = np.array([-50, 700, 320, 320, 639.9], np.float32)
    q["v"][:5] = np.array([240, 240, -60, 520, 479.9], np.float32)
    qd = np.stack([flip_bits(rng, d[s], int(rng.choice([0, 5, 20, 45, 55, 90, 110]))) for s in src])
    return kf, q, qd


def fuse_py(variant, kf, q, qd):
    bi = np.full(len(q), -1, np.int32)
    bd = np.full(len(q), 256, np.int32)
    F = kf.fd
    for i, Q in enumerate(q):
        u, v, ur, pl = f32(Q["u"]), f32(Q["v"]), f32(Q["u_right"]), int(Q["level_hi"])
        best, bidx = 256, -1
        for idx in _features_in_area_py(F, u, v, Q["radius"], -1, -1):
            lv = int(F.kps["octave"][idx])
            if lv < pl - 1 or lv > pl:
                continue
            if variant == 0:
                ex, ey = f32(u - f32(F.kps["x"][idx])), f32(v - f32(F.kps["y"][idx]))
                e2 = f32(f32(ex * ex) + f32(ey * ey))
                lim = 5.99
                if F.u_right[idx] >= 0:
                    er = f32(ur - F.u_right[idx])
                    e2 = f32(e2 + f32(er * er))
                    lim = 7.8
                if float(f32(e2 * kf.is2[lv])) > lim:
                    continue
            dd = _dist(qd[i], F.desc[idx])
            if dd < best:
                best, bidx = dd, idx
        bd[i] = best
        if best <= (100 if variant == 2 else 50):
            bi[i] = bidx
    return bi, bd


def test_fuse_vs_python():
    for variant in (0, 1, 2):
        kf, q, qd = make_fuse_case(10 + variant, n=500, nq=160)
        bi, bd = O.fuse(variant, kf, q, qd)
        pi, pd = fuse_py(variant, kf, q, qd)
        assert np.array_equal(bi, pi) and np.array_equal(bd, pd)
        assert (bi >= 0).sum() > 20 and (bi < 0).sum() > 5


def make_sim3_case(seed):
    kf1, q21, d21 = make_fuse_case(seed, n=600, nq=260)       # queries21 are searched in KF1
    kf2, q12, d12 = make_fuse_case(seed + 50, n=640, nq=280)  # queries12 in KF2
    rng = np.random.default_rng(seed + 99)
    # make the two directions agree for a good share: query a of 12 comes from feature i1 and targets feature j of KF2;
    # add the reverse query (feature j of KF2 -> i1 of KF1)
    index1 = rng.permutation(kf1.n)[:len(q12)].astype(np.int32)
    index2 = rng.permutation(kf2.n)[:len(q21)].astype(np.int32)
    b12, _ = O.fuse(2, kf2, q12, d12)
    k1 = kf1.fd.kps
    for t, (a, j) in enumerate([(a, j) for a, j in enumerate(b12) if j >= 0][:len(q21) // 2]):
        i1 = int(index1[a])
        index2[t] = j
        q21["u"][t], q21["v"][t] = k1["x"][i1], k1["y"][i1]
        lv = int(k1["octave"][i1])
        q21["level_lo"][t], q21["level_hi"][t] = lv - 1, lv
        q21["radius"][t] = f32(3.0) * SCALE[lv]
        d21[t] = kf1.fd.desc[i1]
    # index2 must stay a set of distinct features
    seen, free = set(), [i for i in range(kf2.n) if i not in set(index2.tolist())]
    for t in range(len(index2)):
        if int(index2[t]) in seen:
            index2[t] = free.pop()
        seen.add(int(index2[t]))
    return kf1, kf2, q12, d12, index1, q21, d21, index2


def test_search_by_sim3_vs_python():
    kf1, kf2, q12, d12, index1, q21, d21, index2 = make_sim3_case(21)
    nf, m12 = O.search_by_sim3(kf1, kf2, q12, d12, index1, q21, d21, index2)
    b12, _ = fuse_py(2, kf2, q12, d12)
    b21, _ = fuse_py(2, kf1, q21, d21)
    vn1 = np.full(kf1.n, -1); vn2 = np.full(kf2.n, -1)
    vn1[index1] = b12
    vn2[index2] = b21
    exp = np.array([vn1[i] if vn1[i] >= 0 and vn2[vn1[i]] == i else -1 for i in range(kf1.n)], np.int32)
    assert np.array_equal(m12, exp) and nf == int((exp >= 0).sum())
    assert nf > 20 and (vn1 >= 0).sum() > nf      # some one-directional matches are rejected by the agreement check
