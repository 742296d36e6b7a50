"""Seeded matcher cases in REFERENCE terms (frames, map points with the fields Tracking / LocalMapping leave on them),
each with two runners:

    run_reference(case)  the reference's own code (oracle/_ref/libref_matchers.so through tests/ref2_lib.py)
    run_oracle(case)     the oracle (oracle/liborb_oracle.so) on the flat arrays the C ABI takes; the flattening below is
                         the test-side twin of the host wrapper's (orb_slam2_map_b200/host/orbx.hpp)

Both return a dict of numpy arrays that must be identical.  tests/golden/make_golden_matchers.py freezes the reference's
outputs into tests/golden/ref_matchers.npz so the oracle stays pinned where the reference library is absent.
TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import functools

import numpy as np

import oracle_lib as O
from orb_slam2_map_b200 import synth

F32 = np.float32
TH_HIGH, TH_LOW = 100, 50


@functools.lru_cache(maxsize=None)
def _extracted(w, h, nf, seed, kind="canvas"):
    if kind == "canvas":
        img = synth.textured_canvas(w, h, seed=seed)
    else:
        img = synth.frame_sequence(seed + 2, w, h, seed=3)[seed + 1]
    ex = O.Extractor(nfeatures=nf)
    k, d = ex.extract(img)
    return k, d, ex.tables()["scale"].copy()


def _noisy(desc, r, flips=12, p=0.6):
    qd = desc.copy()
    n = len(qd)
    pos = r.integers(0, 256, (n, flips))
    for j in range(flips):
        use = r.random(n) < p
        qd[use, pos[use, j] // 8] ^= (1 << (pos[use, j] % 8)).astype(np.uint8)
    return qd


def _frame(case, key="frame", B=None):
    f = case[key]
    return (B or O).FrameData(f["kps"], f["desc"], f["w"], f["h"], u_right=f.get("u_right"), scale_factors=f.get("scale"),
                       bounds=f.get("bounds"))


# ------------------------------------------------------------------------------------------------------------------
# DescriptorDistance
# ------------------------------------------------------------------------------------------------------------------
def case_distance(seed=0):
    r = np.random.default_rng(seed)
    a = r.integers(0, 256, (260, 32), dtype=np.uint8)
    b = r.integers(0, 256, (260, 32), dtype=np.uint8)
    a[0] = 0; b[0] = 0
    a[1] = 0; b[1] = 255
    a[2] = 0; b[2] = 0; b[2, 17] = 0x10
    for i in range(3, 40):                       # every byte position once
        b[i] = a[i]; b[i, i - 3 if i - 3 < 32 else 0] ^= 0x81
    return dict(kind="distance", a=a, b=b)


# ------------------------------------------------------------------------------------------------------------------
# Frame::GetFeaturesInArea (+ AssignFeaturesToGrid / PosInGrid)
# ------------------------------------------------------------------------------------------------------------------
def case_area(seed=0, shifted=False):
    k, d, sf = _extracted(640, 480, 1000, 5 + seed)
    r = np.random.default_rng(100 + seed)
    k = k.copy()
    bounds = None
    if shifted:
        # undistortion pushes keypoints outside the image and the bounds are not the image: the "last half cell is not
        # indexed" quirk (PosInGrid rounds, src/Frame.cc:382-392) and negative coordinates
        k["x"] = (k["x"] * F32(1.04) - F32(12.5)).astype(F32)
        k["y"] = (k["y"] * F32(1.03) - F32(7.25)).astype(F32)
        bounds = (-10.5, 655.25, -6.0, 489.5)
    nq = 500
    q = np.zeros((nq, 5), F32)
    q[:, 0] = r.uniform(-40, 700, nq)
    q[:, 1] = r.uniform(-40, 520, nq)
    q[:, 2] = r.choice([3.0, 10.0, 25.5, 100.0, 250.0], nq)
    lv = [(-1, -1), (0, 0), (0, -1), (-1, 3), (2, -1), (1, 2), (3, 3), (0, 7), (7, 7), (1, -1), (-1, 0), (5, 2)]
    pick = r.integers(0, len(lv), nq)
    q[:, 3] = [lv[i][0] for i in pick]
    q[:, 4] = [lv[i][1] for i in pick]
    # exact cell borders
    q[:20, 0] = np.arange(20) * 10.0
    q[:20, 1] = np.arange(20) * 10.0
    return dict(kind="area", frame=dict(kps=k, desc=d, w=640, h=480, scale=sf, bounds=bounds), queries=q)


# ------------------------------------------------------------------------------------------------------------------
# SearchForInitialization
# ------------------------------------------------------------------------------------------------------------------
def case_search_init(seed=0, window=100, nnratio=0.9, ori=True, collide=False):
    k1, d1, sf = _extracted(640, 480, 1000, seed, kind="seq")
    k2, d2, _ = _extracted(640, 480, 1000, seed + 1, kind="seq")
    d1, d2 = d1.copy(), d2.copy()
    if collide:
        # many F1 descriptors equal -> several queries want the same F2 keypoint (steal-back, vMatchedDistance skips),
        # and duplicated F2 descriptors make best == second best
        r = np.random.default_rng(seed)
        lvl0 = np.nonzero(k1["octave"] == 0)[0]
        d1[lvl0[::3]] = d1[lvl0[0]]
        t0 = np.nonzero(k2["octave"] == 0)[0]
        d2[t0[::4]] = d2[t0[1]]
        d2[t0[5]] = d1[lvl0[0]]
    prev = np.stack([k1["x"], k1["y"]], axis=1).astype(F32)
    return dict(kind="search_init", f1=dict(kps=k1, desc=d1, w=640, h=480, scale=sf), f2=dict(kps=k2, desc=d2, w=640, h=480, scale=sf),
                prev=prev, window=window, nnratio=nnratio, ori=ori)


# ------------------------------------------------------------------------------------------------------------------
# Frame::ComputeStereoMatches
# ------------------------------------------------------------------------------------------------------------------
@functools.lru_cache(maxsize=None)
def _stereo(w, h, nf, seed):
    left, right = synth.stereo_pair(w, h, seed=seed)
    el, er = O.Extractor(nf), O.Extractor(nf)
    kl, dl = el.extract(left)
    kr, dr = er.extract(right)
    pl = [el.level(l) for l in range(8)]
    pr = [er.level(l) for l in range(8)]
    t = el.tables()
    return kl, dl, kr, dr, pl, pr, t["scale"].copy(), t["inv_scale"].copy()


def case_stereo(w=640, h=240, nf=800, seed=2, bf=386.1448, fx=718.856, left_frac=1.0, right_frac=1.0, swap=False):
    """left_frac / right_frac keep a prefix of an eye's keypoints (eyes with different keypoint counts); swap: the eyes the other
    way round (disparities of the wrong sign: almost nothing matches, the search structure is walked all the same)"""
    kl, dl, kr, dr, pl, pr, sc, isc = _stereo(w, h, nf, seed)
    nl, nr = int(len(kl) * left_frac), int(len(kr) * right_frac)
    kl, dl, kr, dr = kl[:nl], dl[:nl], kr[:nr], dr[:nr]
    if swap:
        kl, dl, kr, dr, pl, pr = kr, dr, kl, dl, pr, pl
    return dict(kind="stereo", kl=kl, dl=dl, kr=kr, dr=dr, pl=pl, pr=pr, scale=sc, inv_scale=isc, bf=bf, fx=fx)


# ------------------------------------------------------------------------------------------------------------------
# SearchByProjection x4.  A frame + "map points" that project near its own keypoints.
# ------------------------------------------------------------------------------------------------------------------
def _proj_base(seed, nq, dup):
    k, d, sf = _extracted(640, 480, 1000, 20 + seed)
    r = np.random.default_rng(300 + seed)
    n = len(k)
    src = r.integers(0, n, nq)
    u = (k["x"][src] + r.normal(0, 3, nq)).astype(F32)
    v = (k["y"][src] + r.normal(0, 3, nq)).astype(F32)
    # a few outside the image bounds (the reference skips them)
    u[:6] = [-3.0, 641.5, 100.0, 200.0, 0.0, 640.0]
    v[:6] = [50.0, 60.0, -0.5, 480.25, 0.0, 480.0]
    qd = _noisy(d[src], r)
    if dup:
        qd[::3] = qd[0]
    level = k["octave"][src].astype(np.int32)
    angle = ((k["angle"][src] + r.normal(0, 8, nq)) % 360).astype(F32)
    u_right = np.where(r.random(n) < 0.6, k["x"] - 20.0 + r.normal(0, 4, n), -1.0).astype(F32)
    slot_state = r.choice([0, 0, 0, 0, 0, 0, 0, 0, 1, 2], n).astype(np.uint8)
    nobs = r.choice([0, 1, 3], nq, p=[0.3, 0.3, 0.4]).astype(np.int32)
    frame = dict(kps=k, desc=d, w=640, h=480, scale=sf, u_right=u_right)
    return frame, r, src, u, v, qd, level, angle, slot_state, nobs


def case_proj_localmap(seed=0, th=1.0, nnratio=0.8, dup=False, nq=900):
    frame, r, src, u, v, qd, level, angle, slot_state, nobs = _proj_base(seed, nq, dup)
    view_cos = r.choice([0.9, 0.9985, 0.998, 0.99801, 0.5], nq).astype(F32)
    in_view = (r.random(nq) < 0.9).astype(np.uint8)
    proj_xr = (u - F32(20.0) + r.normal(0, 3, nq)).astype(F32)
    return dict(kind="proj_localmap", frame=frame, slot_state=slot_state, proj_x=u, proj_y=v, proj_xr=proj_xr, level=level,
                view_cos=view_cos, in_view=in_view, qdesc=qd, nobs=nobs, th=th, nnratio=nnratio)


def case_proj_lastframe(seed=0, th=15.0, last_tz=0.0, mono=False, ori=True, dup=False, nq=900, bf=20.0):
    frame, r, src, u, v, qd, level, angle, slot_state, nobs = _proj_base(seed, nq, dup)
    outlier = (r.random(nq) < 0.05).astype(np.uint8)
    return dict(kind="proj_lastframe", frame=frame, slot_state=slot_state, bf=bf, u=u, v=v, octave=level, angle=angle, qdesc=qd,
                nobs=nobs, outlier=outlier, last_tz=last_tz, th=th, mono=mono, ori=ori)


def _max_dist_for_level(u, v, level, sf):
    """mfMaxDistance that makes MapPoint::PredictScale(dist, log(1.2)) = ceil(log(max/dist)/log(sf)) return `level`"""
    dist = np.sqrt(u.astype(np.float64) ** 2 + v.astype(np.float64) ** 2 + 1.0).astype(F32)
    return (dist.astype(np.float64) * float(sf[1]) ** (level - 0.5)).astype(F32)


def case_proj_keyframe(seed=0, th=10.0, orb_dist=64, ori=True, dup=False, nq=900):
    frame, r, src, u, v, qd, level, angle, slot_state, nobs = _proj_base(seed, nq, dup)
    max_dist = _max_dist_for_level(u, v, level, frame["scale"])
    found = (r.random(nq) < 0.08).astype(np.uint8)
    return dict(kind="proj_keyframe", frame=frame, slot_state=slot_state, u=u, v=v, level=level, max_dist=max_dist, angle=angle,
                qdesc=qd, already_found=found, th=th, orb_dist=orb_dist, ori=ori)


def case_proj_sim3(seed=0, th=10, dup=False, nq=900):
    frame, r, src, u, v, qd, level, angle, slot_state, nobs = _proj_base(seed, nq, dup)
    frame = dict(frame); frame["u_right"] = None
    level = np.maximum(level, 0)
    max_dist = _max_dist_for_level(u, v, level, frame["scale"])
    return dict(kind="proj_sim3", frame=frame, matched_state=(slot_state != 0).astype(np.uint8), u=u, v=v, level=level,
                max_dist=max_dist, qdesc=qd, th=th)


# ------------------------------------------------------------------------------------------------------------------
# SURVEY §8(f) rows
# ------------------------------------------------------------------------------------------------------------------
def case_bow_transform(seed=0, k=6, depth=4, n=600):
    voc = O.make_vocabulary(k=k, depth=depth, seed=seed)
    r = np.random.default_rng(400 + seed)
    pick = r.integers(1, voc["nnodes"], n)
    feats = _noisy(voc["desc"][pick], r, flips=30, p=0.5)
    feats[:8] = voc["desc"][pick[:8]]                      # exact node descriptors: distance ties between siblings' subtrees
    return dict(kind="bow_transform", voc=voc, feats=feats)


def _bow_sides(seed, voc, levelsup):
    k1, d1, sf = _extracted(640, 480, 1000, seed, kind="seq")
    k2, d2, _ = _extracted(640, 480, 1000, seed + 1, kind="seq")
    r = np.random.default_rng(500 + seed)
    out = []
    for k, d in ((k1, d1), (k2, d2)):
        _, w, nid = O.bow_transform(voc, d, levelsup)
        flags = (w > 0).astype(np.uint8)                   # bit 0: the feature's word has a weight -> it is in mFeatVec
        has = r.random(len(k)) < 0.65
        bad = has & (r.random(len(k)) < 0.06)
        flags |= (has & ~bad).astype(np.uint8) << 1
        flags |= bad.astype(np.uint8) << 2
        out.append(dict(kps=k, desc=d, w=640, h=480, scale=sf, node=nid, flags=flags))
    return out


def case_search_by_bow(variant, seed=0, nnratio=0.7, ori=True, levelsup=2, dup=False):
    voc = O.make_vocabulary(k=5, depth=4, seed=7)
    a, b = _bow_sides(seed, voc, levelsup)
    if dup:
        a = dict(a); b = dict(b)
        a["desc"] = a["desc"].copy(); b["desc"] = b["desc"].copy()
        a["desc"][::5] = a["desc"][0]; b["desc"][::7] = a["desc"][0]
    return dict(kind="search_by_bow", variant=variant, a=a, b=b, nnratio=nnratio, ori=ori)


def case_triangulation(seed=0, only_stereo=False, ori=True):
    voc = O.make_vocabulary(k=5, depth=4, seed=7)
    a, b = _bow_sides(seed, voc, 2)
    r = np.random.default_rng(600 + seed)
    for side in (a, b):
        n = len(side["kps"])
        side["u_right"] = np.where(r.random(n) < 0.5, side["kps"]["x"] - 20.0, -1.0).astype(F32)
    # the synthetic sequence translates by (1, 0.5) px per frame: F12 = [t]x, epipolar lines parallel to t
    f12 = np.array([[0, 0, 0.5], [0, 0, -1.0], [-0.5, 1.0, 0]], F32)
    return dict(kind="triangulation", a=a, b=b, f12=f12, epipole=(F32(320.5), F32(240.25)), only_stereo=only_stereo, ori=ori)


def case_fuse(variant, seed=0, th=3.0, dup=False, nq=900, bf=40.0):
    frame, r, src, u, v, qd, level, angle, slot_state, nobs = _proj_base(seed, nq, dup)
    level = np.maximum(level, 0)
    max_dist = _max_dist_for_level(u, v, level, frame["scale"])
    return dict(kind="fuse", variant=variant, frame=frame, bf=bf, u=u, v=v, level=level, max_dist=max_dist, qdesc=qd,
                nobs=np.maximum(nobs, 1), th=th)


def case_sim3(seed=0, th=7.5):
    k1, d1, sf = _extracted(640, 480, 1000, seed, kind="seq")
    k2, d2, _ = _extracted(640, 480, 1000, seed + 1, kind="seq")
    r = np.random.default_rng(700 + seed)
    sides = []
    for k in (k1, k2):
        n = len(k)
        has = (r.random(n) < 0.8).astype(np.uint8)
        u = (k["x"] + r.normal(0, 1.5, n)).astype(F32)
        v = (k["y"] + r.normal(0, 1.5, n)).astype(F32)
        lv = k["octave"].astype(np.int32)
        sides.append((has, u, v, lv, _max_dist_for_level(u, v, lv, sf)))
    pre12 = np.full(len(k1), -1, np.int32)
    cand = np.nonzero(sides[0][0])[0][::15]
    tgt = np.nonzero(sides[1][0])[0]
    pre12[cand] = r.choice(tgt, len(cand), replace=False)
    return dict(kind="sim3", f1=dict(kps=k1, desc=d1, w=640, h=480, scale=sf), f2=dict(kps=k2, desc=d2, w=640, h=480, scale=sf),
                s1=sides[0], s2=sides[1], pre12=pre12, th=th)


def case_distinctive(seed=0):
    r = np.random.default_rng(800 + seed)
    groups = []
    for n in [1, 2, 3, 4, 5, 8, 13, 21, 40, 64]:
        base = r.integers(0, 256, 32, dtype=np.uint8)
        g = _noisy(np.repeat(base[None], n, 0), r, flips=40, p=0.5)
        groups.append(g)
    groups.append(np.repeat(groups[3][:1], 6, 0))                 # all equal: every median 0, first wins
    g = groups[5].copy(); g[3] = g[1]; groups.append(g)           # duplicated rows: ties between medians
    return dict(kind="distinctive", groups=groups)


def _queries(n):
    q = np.zeros(n, O.PROJ_DTYPE)
    q["level_lo"] = -1
    q["level_hi"] = -1
    q["max_right_err"] = -1.0
    return q


def _remap(assigned, keep):
    """oracle query indices refer to the flattened (kept) list: map them back to the case's map point indices"""
    idx = np.nonzero(keep)[0]
    out = np.full_like(assigned, -1)
    m = assigned >= 0
    out[m] = idx[assigned[m]]
    return out


# ------------------------------------------------------------------------------------------------------------------
def run_reference(c):
    import ref2_lib as R
    k = c["kind"]
    if k == "distance":
        return dict(dist=np.array([R.descriptor_distance(a, b) for a, b in zip(c["a"], c["b"])], np.int32))
    if k == "area":
        f = _frame(c)
        res = [R.features_in_area(f, *map(float, q[:3]), int(q[3]), int(q[4])) for q in c["queries"]]
        return dict(counts=np.array([len(x) for x in res], np.int32), idx=np.concatenate(res + [np.zeros(0, np.int32)]).astype(np.int32))
    if k == "search_init":
        n, m12, pm = R.search_for_initialization(_frame(c, "f1"), _frame(c, "f2"), c["prev"], c["window"], c["nnratio"], c["ori"])
        return dict(n=np.int32(n), m12=m12, prev=pm)
    if k == "stereo":
        ur, dp = R.compute_stereo_matches(c["kl"], c["dl"], c["kr"], c["dr"], c["pl"], c["pr"], c["scale"], c["inv_scale"], c["bf"], c["fx"])
        return dict(u_right=ur, depth=dp)
    if k == "proj_localmap":
        n, a = R.proj_localmap(_frame(c), c["slot_state"], c["proj_x"], c["proj_y"], c["proj_xr"], c["level"], c["view_cos"],
                               c["in_view"], c["qdesc"], c["nobs"], c["th"], c["nnratio"], False)
        return dict(n=np.int32(n), assigned=np.where(a == -2, -1, a).astype(np.int32))
    if k == "proj_lastframe":
        n, a = R.proj_lastframe(_frame(c), c["slot_state"], c["bf"], c["u"], c["v"], c["octave"], c["angle"], c["qdesc"], c["nobs"],
                                c["outlier"], c["last_tz"], c["th"], c["mono"], 0.9, c["ori"])
        return dict(n=np.int32(n), assigned=np.where(a == -2, -1, a).astype(np.int32))
    if k == "proj_keyframe":
        n, a, lv = R.proj_keyframe(_frame(c), c["slot_state"], c["u"], c["v"], c["max_dist"], c["angle"], c["qdesc"], c["already_found"],
                                   c["th"], c["orb_dist"], 0.9, c["ori"])
        assert np.array_equal(lv, c["level"]), "MapPoint::PredictScale did not return the level the case asked for"
        return dict(n=np.int32(n), assigned=np.where(a == -2, -1, a).astype(np.int32))
    if k == "proj_sim3":
        n, a, lv = R.proj_sim3(_frame(c), c["matched_state"], c["u"], c["v"], c["max_dist"], c["qdesc"], c["th"])
        assert np.array_equal(lv, c["level"])
        return dict(n=np.int32(n), assigned=np.where(a == -2, -1, a).astype(np.int32))
    if k == "bow_transform":
        import ref2_lib
        voc = ref2_lib.RefVocabulary(c["voc"])
        out = {}
        for ls in range(0, c["voc"]["depth"] + 2):
            w, val, nid = voc.transform(c["feats"], ls)
            bw, bv, fn, ff = voc.transform_all(c["feats"], ls)
            out.update({f"word{ls}": w, f"weight{ls}": val, f"node{ls}": nid, f"bow_word{ls}": bw, f"bow_value{ls}": bv,
                        f"fv_node{ls}": fn, f"fv_feat{ls}": ff})
        return out
    if k == "search_by_bow":
        a, b = c["a"], c["b"]
        n, ma, mb = R.search_by_bow(c["variant"], _frame(c, "a"), a["node"], a["flags"], _frame(c, "b"), b["node"], b["flags"], c["nnratio"], c["ori"])
        return dict(n=np.int32(n), match_a=ma, match_b=mb)
    if k == "triangulation":
        a, b = c["a"], c["b"]
        n, m = R.search_for_triangulation(_frame(c, "a"), a["node"], a["flags"], _frame(c, "b"), b["node"], b["flags"], c["f12"], c["epipole"],
                                          c["only_stereo"], c["ori"])
        return dict(n=np.int32(n), matches12=m)
    if k == "fuse":
        n, bi, lv = R.fuse(c["variant"], _frame(c), c["bf"], c["u"], c["v"], c["max_dist"], c["qdesc"], c["nobs"], c["th"])
        assert np.array_equal(lv, c["level"])
        return dict(n=np.int32(n), best_idx=bi)
    if k == "sim3":
        (h1, u1, v1, l1, m1), (h2, u2, v2, l2, m2) = c["s1"], c["s2"]
        n, m, lv1, lv2 = R.search_by_sim3(_frame(c, "f1"), _frame(c, "f2"), h1, u1, v1, m1, h2, u2, v2, m2, c["pre12"], c["th"])
        assert np.array_equal(lv1[h1 > 0], l1[h1 > 0]) and np.array_equal(lv2[h2 > 0], l2[h2 > 0])
        return dict(n=np.int32(n), match12=m)
    if k == "distinctive":
        return dict(desc=np.stack([R.distinctive(g) for g in c["groups"]]))
    raise KeyError(k)


def run_oracle(c, B=None):
    """The case through the flat interface of include/orbx.h.  B = the implementation behind it: the CPU oracle (default) or an
    object with the same functions over the CUDA library (tests/test_gpu_vs_reference_golden.py)."""
    B = B or O
    k = c["kind"]
    if k == "distance":
        return dict(dist=np.array([B.descriptor_distance(a, b) for a, b in zip(c["a"], c["b"])], np.int32))
    if k == "area":
        f = _frame(c, B=B)
        res = [B.features_in_area(f, *map(float, q[:3]), int(q[3]), int(q[4])) for q in c["queries"]]
        return dict(counts=np.array([len(x) for x in res], np.int32), idx=np.concatenate(res + [np.zeros(0, np.int32)]).astype(np.int32))
    if k == "search_init":
        n, m12, pm, _ = B.search_for_initialization(_frame(c, "f1", B=B), _frame(c, "f2", B=B), c["prev"], c["window"], c["nnratio"], c["ori"])
        return dict(n=np.int32(n), m12=m12, prev=pm)
    if k == "stereo":
        ur, dp, _ = B.compute_stereo_matches(c["kl"], c["dl"], c["kr"], c["dr"], c["pl"], c["pr"], c["scale"], c["inv_scale"], c["bf"], c["fx"])
        return dict(u_right=ur, depth=dp)
    if k == "bow_transform":
        out = {}
        for ls in range(0, c["voc"]["depth"] + 2):
            w, val, nid = B.bow_transform(c["voc"], c["feats"], ls)
            # TemplatedVocabulary::transform (TemplatedVocabulary.h:1140-1218), TF_IDF + L1: BowVector[word] += weight and
            # FeatureVector[node].push_back(i) for the features whose word has a weight, then the L1 normalisation
            use = val > 0
            words = np.unique(w[use])
            vals = np.zeros(len(words), np.float64)
            for i in np.nonzero(use)[0]:
                vals[np.searchsorted(words, w[i])] += val[i]
            norm = 0.0
            for x in vals:
                norm += abs(x)
            if norm > 0.0:
                vals = vals / norm
            order = np.lexsort((np.nonzero(use)[0], nid[use]))
            out.update({f"word{ls}": w, f"weight{ls}": val, f"node{ls}": nid, f"bow_word{ls}": words.astype(np.uint32), f"bow_value{ls}": vals,
                        f"fv_node{ls}": nid[use][order].astype(np.uint32), f"fv_feat{ls}": np.nonzero(use)[0][order].astype(np.uint32)})
        return out
    if k == "search_by_bow":
        a, b = c["a"], c["b"]
        fl = lambda x: (x["flags"] & 3).astype(np.uint8)          # flat form: bit 1 = holds a map point that is NOT bad
        n, ma, mb = B.search_by_bow(c["variant"], a["desc"], a["kps"]["angle"], a["node"], fl(a), b["desc"], b["kps"]["angle"], b["node"], fl(b),
                                    c["nnratio"], c["ori"])
        return dict(n=np.int32(n), match_a=ma, match_b=mb)
    if k == "triangulation":
        a, b = c["a"], c["b"]
        # KeyFrame::GetMapPoint(idx) != NULL is all the reference asks (bad points included)
        fl = lambda x: ((x["flags"] & 1) | (((x["flags"] & 6) != 0).astype(np.uint8) << 1)).astype(np.uint8)
        mk = lambda x: B.KeyFrameData(x["kps"], x["desc"], 640, 480, u_right=x["u_right"], scale_factors=x["scale"],
                                      level_sigma2=(x["scale"] * x["scale"]).astype(F32), node_id=x["node"], flags=fl(x))
        n, m = B.search_for_triangulation(mk(a), mk(b), c["f12"], c["epipole"], c["only_stereo"], c["ori"])
        return dict(n=np.int32(n), matches12=m)
    if k == "fuse":
        fr = c["frame"]
        sf = fr["scale"]
        kf = B.KeyFrameData(fr["kps"], fr["desc"], 640, 480, u_right=fr["u_right"], scale_factors=sf, level_sigma2=(sf * sf).astype(F32))
        u, v = c["u"], c["v"]
        keep = (u >= 0) & (u < 640) & (v >= 0) & (v < 480)          # KeyFrame::IsInImage
        q = _queries(int(keep.sum()))
        lv = c["level"][keep]
        q["u"] = u[keep]; q["v"] = v[keep]
        q["radius"] = (F32(c["th"]) * sf[lv]).astype(F32)
        q["u_right"] = (u[keep] - F32(c["bf"])).astype(F32)
        q["level_lo"] = lv - 1; q["level_hi"] = lv
        bi, bd = B.fuse(c["variant"], kf, q, c["qdesc"][keep])
        out = np.full(len(u), -1, np.int32)
        out[keep] = bi
        return dict(n=np.int32((bi >= 0).sum()), best_idx=out)
    if k == "sim3":
        (h1, u1, v1, l1, m1), (h2, u2, v2, l2, m2) = c["s1"], c["s2"]
        f1, f2 = c["f1"], c["f2"]
        sf = f1["scale"]
        mk = lambda x: B.KeyFrameData(x["kps"], x["desc"], 640, 480, scale_factors=sf, level_sigma2=(sf * sf).astype(F32))
        pre = c["pre12"]
        done2 = np.zeros(len(h2), bool); done2[pre[pre >= 0]] = True
        def side(has, u, v, lv, skip):
            keep = (has > 0) & ~skip & (u >= 0) & (u < 640) & (v >= 0) & (v < 480)
            idx = np.nonzero(keep)[0].astype(np.int32)
            q = _queries(len(idx))
            q["u"] = u[idx]; q["v"] = v[idx]
            q["radius"] = (F32(c["th"]) * sf[lv[idx]]).astype(F32)
            q["level_lo"] = lv[idx] - 1; q["level_hi"] = lv[idx]
            return q, idx
        q12, i1 = side(h1, u1, v1, l1, pre >= 0)
        q21, i2 = side(h2, u2, v2, l2, done2)
        n, m = B.search_by_sim3(mk(f1), mk(f2), q12, f1["desc"][i1], i1, q21, f2["desc"][i2], i2)
        return dict(n=np.int32(n), match12=m)
    if k == "distinctive":
        off = np.cumsum([0] + [len(g) for g in c["groups"]]).astype(np.int32)
        alld = np.concatenate(c["groups"])
        idx, _ = B.distinctive_descriptors(alld, off)
        return dict(desc=np.stack([g[i] for g, i in zip(c["groups"], idx)]))
    f = _frame(c, B=B)
    sf = c["frame"]["scale"]
    b = c["frame"].get("bounds") or (0.0, float(c["frame"]["w"]), 0.0, float(c["frame"]["h"]))
    if k == "proj_localmap":
        # src/ORBmatcher.cc:45-77: in-view points only; r = RadiusByViewingCos(cos) [* th]; window r * scale[level]
        keep = c["in_view"].astype(bool)
        q = _queries(int(keep.sum()))
        # RadiusByViewingCos compares the float cosine with the DOUBLE literal 0.998 (src/ORBmatcher.cc:131-137)
        rr = np.where(c["view_cos"].astype(np.float64) > 0.998, F32(2.5), F32(4.0)).astype(F32)
        if c["th"] != 1.0:
            rr = (rr * F32(c["th"])).astype(F32)
        rad = (rr * sf[c["level"]]).astype(F32)[keep]
        q["u"] = c["proj_x"][keep]; q["v"] = c["proj_y"][keep]; q["radius"] = rad
        q["min_level"] = c["level"][keep] - 1; q["max_level"] = c["level"][keep]
        q["u_right"] = c["proj_xr"][keep]; q["max_right_err"] = rad
        n, a, _, _, _ = B.search_by_projection(0, f, q, c["qdesc"][keep], (c["nobs"][keep] > 0).astype(np.uint8),
                                               (c["slot_state"] == 1).astype(np.uint8), TH_HIGH, c["nnratio"], False)
        return dict(n=np.int32(n), assigned=_remap(a, keep))
    if k == "proj_lastframe":
        # src/ORBmatcher.cc:1328-1392: outliers and points outside the bounds are skipped; window th * scale[octave];
        # level window by the motion direction
        u, v = c["u"], c["v"]
        keep = (c["outlier"] == 0) & ~((u < b[0]) | (u > b[1])) & ~((v < b[2]) | (v > b[3]))
        mb = F32(c["bf"]) / F32(1.0)
        fwd = (F32(c["last_tz"]) > mb) and not c["mono"]
        bwd = (-F32(c["last_tz"]) > mb) and not c["mono"]
        q = _queries(int(keep.sum()))
        octv = c["octave"][keep]
        q["u"] = u[keep]; q["v"] = v[keep]
        q["radius"] = (F32(c["th"]) * sf[octv]).astype(F32)
        if fwd:
            q["min_level"] = octv; q["max_level"] = -1
        elif bwd:
            q["min_level"] = 0; q["max_level"] = octv
        else:
            q["min_level"] = octv - 1; q["max_level"] = octv + 1
        q["u_right"] = (u[keep] - F32(c["bf"])).astype(F32)           # u - mbf * invzc, invzc = 1
        q["max_right_err"] = q["radius"]
        q["angle"] = c["angle"][keep]
        n, a, _, _, _ = B.search_by_projection(1, f, q, c["qdesc"][keep], (c["nobs"][keep] > 0).astype(np.uint8),
                                               (c["slot_state"] == 1).astype(np.uint8), TH_HIGH, 0.9, c["ori"])
        return dict(n=np.int32(n), assigned=_remap(a, keep))
    if k == "proj_keyframe":
        # src/ORBmatcher.cc:1472-1530: already-found and out-of-bounds points skipped; any non-NULL slot blocks
        u, v = c["u"], c["v"]
        keep = (c["already_found"] == 0) & ~((u < b[0]) | (u > b[1])) & ~((v < b[2]) | (v > b[3]))
        q = _queries(int(keep.sum()))
        lv = c["level"][keep]
        q["u"] = u[keep]; q["v"] = v[keep]
        q["radius"] = (F32(c["th"]) * sf[lv]).astype(F32)
        q["min_level"] = lv - 1; q["max_level"] = lv + 1
        q["angle"] = c["angle"][keep]
        n, a, _, _, _ = B.search_by_projection(2, f, q, c["qdesc"][keep], None, (c["slot_state"] != 0).astype(np.uint8),
                                               c["orb_dist"], 0.9, c["ori"])
        return dict(n=np.int32(n), assigned=_remap(a, keep))
    if k == "proj_sim3":
        # src/ORBmatcher.cc:290-403: KeyFrame::IsInImage, KeyFrame::GetFeaturesInArea (no level window), octave filter
        u, v = c["u"], c["v"]
        keep = (u >= b[0]) & (u < b[1]) & (v >= b[2]) & (v < b[3])
        q = _queries(int(keep.sum()))
        lv = c["level"][keep]
        q["u"] = u[keep]; q["v"] = v[keep]
        q["radius"] = (F32(c["th"]) * sf[lv]).astype(F32)
        q["min_level"] = -1; q["max_level"] = -1
        # kpLevel < nPredictedLevel-1 || kpLevel > nPredictedLevel (:374-377); a negative bound switches the filter off in
        # the flat form, so level 0 is passed as [0, 0]
        q["level_lo"] = np.maximum(lv - 1, 0); q["level_hi"] = lv
        n, a, _, _, _ = B.search_by_projection(3, f, q, c["qdesc"][keep], None, c["matched_state"], TH_LOW, 0.75, False)
        return dict(n=np.int32(n), assigned=_remap(a, keep))
    raise KeyError(k)


# every case the pinning tests run: name -> builder
CASES = {
    "distance": lambda: case_distance(),
    "area": lambda: case_area(0),
    "area_shifted_bounds": lambda: case_area(1, shifted=True),
    "search_init_default": lambda: case_search_init(0, 100, 0.9, True),
    "search_init_w30_noori": lambda: case_search_init(1, 30, 0.7, False),
    "search_init_w250": lambda: case_search_init(2, 250, 0.95, True),
    "search_init_collisions": lambda: case_search_init(0, 100, 0.9, True, collide=True),
    "stereo_640x240": lambda: case_stereo(640, 240, 800, 2),
    "stereo_kitti": lambda: case_stereo(1241, 376, 2000, 5),
    "stereo_fewer_left": lambda: case_stereo(640, 240, 800, 3, left_frac=0.5),
    "stereo_fewer_right_swapped": lambda: case_stereo(640, 240, 800, 4, right_frac=0.34, swap=True),
    "search_init_w10": lambda: case_search_init(3, 10, 0.9, True),
    "proj_localmap": lambda: case_proj_localmap(0, 1.0, 0.8),
    "proj_localmap_th3_dup": lambda: case_proj_localmap(1, 3.0, 0.99, dup=True),
    "proj_lastframe": lambda: case_proj_lastframe(0, 15.0, 0.0),
    "proj_lastframe_forward": lambda: case_proj_lastframe(1, 7.0, 50.0, ori=False),
    "proj_lastframe_backward_dup": lambda: case_proj_lastframe(2, 15.0, -50.0, dup=True),
    "proj_lastframe_mono": lambda: case_proj_lastframe(0, 15.0, 50.0, mono=True),
    "proj_keyframe": lambda: case_proj_keyframe(0, 10.0, 64),
    "proj_keyframe_dup_noori": lambda: case_proj_keyframe(1, 15.0, 100, ori=False, dup=True),
    "proj_sim3": lambda: case_proj_sim3(0, 10),
    "proj_sim3_dup": lambda: case_proj_sim3(1, 15, dup=True),
    "bow_transform": lambda: case_bow_transform(0),
    "bow_transform_k9": lambda: case_bow_transform(1, k=9, depth=3),
    "search_by_bow_kf_frame": lambda: case_search_by_bow(0, 0, 0.7, True),
    "search_by_bow_kf_frame_dup_noori": lambda: case_search_by_bow(0, 1, 0.9, False, dup=True),
    "search_by_bow_kf_kf": lambda: case_search_by_bow(1, 0, 0.75, True),
    "search_by_bow_kf_kf_dup": lambda: case_search_by_bow(1, 2, 0.8, True, levelsup=3, dup=True),
    "triangulation": lambda: case_triangulation(0),
    "triangulation_only_stereo_noori": lambda: case_triangulation(1, only_stereo=True, ori=False),
    "fuse_mappoints": lambda: case_fuse(0, 0, 3.0),
    "fuse_mappoints_dup": lambda: case_fuse(0, 1, 4.0, dup=True),
    "fuse_sim3": lambda: case_fuse(1, 0, 4.0),
    "fuse_sim3_dup": lambda: case_fuse(1, 2, 3.0, dup=True),
    "search_by_sim3": lambda: case_sim3(0),
    "search_by_sim3_th12": lambda: case_sim3(2, 12.0),
    "distinctive": lambda: case_distinctive(0),
    # further seeds of the rows with the most state (first-come ownership, ratio tests, rotation culls)
    "proj_localmap_s5": lambda: case_proj_localmap(5, 1.0, 0.8),
    "proj_lastframe_s6": lambda: case_proj_lastframe(6, 15.0, 0.0),
    "proj_keyframe_s7": lambda: case_proj_keyframe(7, 10.0, 64),
    "proj_sim3_s8": lambda: case_proj_sim3(8, 10),
    "search_init_s9": lambda: case_search_init(9, 100, 0.9, True),
    "search_by_bow_kf_frame_s3": lambda: case_search_by_bow(0, 3, 0.7, True),
    "search_by_bow_kf_kf_s4": lambda: case_search_by_bow(1, 4, 0.75, True),
    "triangulation_s5": lambda: case_triangulation(5),
    "fuse_mappoints_s6": lambda: case_fuse(0, 6, 3.0),
    "fuse_sim3_s7": lambda: case_fuse(1, 7, 4.0),
    "search_by_sim3_s8": lambda: case_sim3(8),
}
