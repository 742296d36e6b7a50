"""oracle/bow_oracle.cpp against independent pure-Python statements of DBoW2's per-feature transform
(TemplatedVocabulary.h:1231-1271) and of the two ORBmatcher::SearchByBoW variants (ORBmatcher.cc:159-288, 522-655)."""
import numpy as np
import pytest

import oracle_lib as O


def _dist(a, b):
    return int(np.unpackbits(np.bitwise_xor(a, b)).sum())


def py_transform(voc, feat, levelsup):
    nid_level = voc["depth"] - levelsup
    nid, final, lvl = 0, 0, 0
    while True:
        lvl += 1
        ch = voc["children"][voc["child_begin"][final]: voc["child_begin"][final] + voc["child_count"][final]]
        final = int(ch[0]); best = _dist(feat, voc["desc"][final])
        for c in ch[1:]:
            d = _dist(feat, voc["desc"][int(c)])
            if d < best:
                best, final = d, int(c)
        if lvl == nid_level:
            nid = final
        if voc["child_count"][final] == 0:
            break
    return int(voc["word"][final]), float(voc["weight"][final]), nid


def py_search_by_bow(variant, dA, aA, nA, fA, dB, aB, nB, fB, nnratio, check_ori):
    fvA, fvB = {}, {}
    for i in range(len(dA)):
        if fA[i] & 1: fvA.setdefault(int(nA[i]), []).append(i)
    for i in range(len(dB)):
        if fB[i] & 1: fvB.setdefault(int(nB[i]), []).append(i)
    mA = -np.ones(len(dA), np.int32); mB = -np.ones(len(dB), np.int32)
    hist = [[] for _ in range(30)]
    n = 0
    for node in sorted(set(fvA) & set(fvB)):           # the lower_bound dance visits exactly the common nodes, ascending
        for ia in fvA[node]:
            if not fA[ia] & 2: continue
            b1, b2, bi = 256, 256, -1
            for ib in fvB[node]:
                if mB[ib] >= 0: continue
                if variant == 1 and not fB[ib] & 2: continue
                d = _dist(dA[ia], dB[ib])
                if d < b1: b2, b1, bi = b1, d, ib
                elif d < b2: b2 = d
            ok = b1 <= 50 if variant == 0 else b1 < 50
            if ok and np.float32(b1) < np.float32(nnratio) * np.float32(b2):
                mA[ia], mB[bi] = bi, ia
                rot = np.float32(aA[ia]) - np.float32(aB[bi])
                if rot < 0: rot = np.float32(rot + np.float32(360.0))
                b = int(np.round(np.float32(rot * np.float32(1.0 / 30))))      # round half away == np.round? values are never .5 exactly here
                if b == 30: b = 0
                hist[b].append(ia); n += 1
    if check_ori:
        sizes = [len(h) for h in hist]
        m1 = m2 = m3 = 0; i1 = i2 = i3 = -1
        for i, s in enumerate(sizes):
            if s > m1: m3, m2, m1, i3, i2, i1 = m2, m1, s, i2, i1, i
            elif s > m2: m3, m2, i3, i2 = m2, s, i2, i
            elif s > m3: m3, i3 = s, i
        if m2 < np.float32(0.1) * np.float32(m1): i2 = i3 = -1
        elif m3 < np.float32(0.1) * np.float32(m1): i3 = -1
        for i in range(30):
            if i in (i1, i2, i3): continue
            for ia in hist[i]:
                mB[mA[ia]] = -1; mA[ia] = -1; n -= 1
    return n, mA, mB


def make_pair(voc, n=300, seed=0, levelsup=1):
    rng = np.random.default_rng(seed)
    leaves = np.flatnonzero(voc["child_count"] == 0)
    base = voc["desc"][rng.choice(leaves, n)]
    def noisy(d, nb):
        bits = np.unpackbits(d, axis=1)
        for r in range(len(bits)):
            bits[r, rng.choice(256, nb, replace=False)] ^= 1
        return np.packbits(bits, axis=1)
    dA = noisy(base, 12); dB = noisy(base[rng.permutation(n)], 14)
    aA = (rng.random(n) * 360).astype(np.float32); aB = ((aA[rng.permutation(n)] + rng.normal(0, 4, n)) % 360).astype(np.float32)
    out = []
    for d in (dA, dB):
        wid, w, nid = O.bow_transform(voc, d, levelsup)
        flags = (w > 0).astype(np.uint8) | ((rng.random(n) < 0.8).astype(np.uint8) << 1)
        out.append((d, nid, flags))
    return (out[0][0], aA, out[0][1], out[0][2]), (out[1][0], aB, out[1][1], out[1][2])


def test_transform_matches_python():
    voc = O.make_vocabulary(k=5, depth=3, seed=1)
    rng = np.random.default_rng(2)
    feats = rng.integers(0, 256, (60, 32), dtype=np.uint8)
    leaves = np.flatnonzero(voc["child_count"] == 0)
    feats[:30] = voc["desc"][rng.choice(leaves, 30)]
    for levelsup in (0, 1, 2, 3, 5):
        wid, w, nid = O.bow_transform(voc, feats, levelsup)
        for i, f in enumerate(feats):
            assert (int(wid[i]), float(w[i]), int(nid[i])) == py_transform(voc, f, levelsup)


@pytest.mark.parametrize("variant", [0, 1])
@pytest.mark.parametrize("check_ori", [True, False])
def test_search_by_bow_matches_python(variant, check_ori):
    voc = O.make_vocabulary(k=6, depth=3, seed=3)
    for seed in range(3):
        A, B = make_pair(voc, n=220, seed=seed, levelsup=2 if seed else 1)
        for ratio in (0.7, 0.9):
            n, mA, mB = O.search_by_bow(variant, *A, *B, nnratio=ratio, check_ori=check_ori)
            n2, mA2, mB2 = py_search_by_bow(variant, *A, *B, ratio, check_ori)
            assert n == n2 and np.array_equal(mA, mA2) and np.array_equal(mB, mB2)
            assert n == int((mA >= 0).sum()) == int((mB >= 0).sum())
            if seed == 0 and ratio == 0.9:
                assert n > 20                      # the synthetic pair really matches


def test_distinctive_descriptors_matches_numpy():
    rng = np.random.default_rng(5)
    sizes = [1, 2, 3, 7, 0, 20, 64, 5]
    offsets = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    base = rng.integers(0, 256, 32, dtype=np.uint8)
    desc = np.repeat(base[None], offsets[-1], 0)
    flips = rng.integers(0, 256, desc.shape, dtype=np.uint8) & rng.integers(0, 256, desc.shape, dtype=np.uint8) & rng.integers(0, 256, desc.shape, dtype=np.uint8)
    desc ^= flips
    idx, med = O.distinctive_descriptors(desc, offsets)
    for p, n in enumerate(sizes):
        if n == 0:
            assert idx[p] == -1; continue
        d = desc[offsets[p]:offsets[p + 1]]
        D = np.array([[_dist(a, b) for b in d] for a in d])
        meds = np.sort(D, axis=1)[:, int(0.5 * (n - 1))]       # MapPoint.cc:291-293
        assert idx[p] == int(np.argmin(meds)) and med[p] == int(meds.min())
