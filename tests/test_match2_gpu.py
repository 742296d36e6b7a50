"""GPU parity (through the C ABI) of SearchByProjection x4, ComputeStereoMatches (stateless and
fused with the stereo extraction) and the brute-force 2-NN search, against the CPU oracle.
Indices, counts, distances: identical.  uRight/depth floats: identical bits (tolerance allowed by
north_star is 1e-5 relative; the kernels reproduce the float sequence exactly)."""
import numpy as np
import pytest
import torch

import oracle_lib as O
from orb_slam2_map_b200 import capi, synth

pytestmark = pytest.mark.gpu


def _frame_with_queries(seed, nq=900, dup=False):
    """An extracted frame + projected-map-point style queries around its own keypoints."""
    img = synth.textured_canvas(640, 480, seed=seed)
    orc = O.Extractor()
    k, d = orc.extract(img)
    return _queries_from_keypoints(k, d, orc.tables()["scale"], seed, nq, dup)


def _queries_from_keypoints(k, d, scale, seed, nq=900, dup=False):
    r = np.random.default_rng(seed)
    n = len(k)
    src = r.integers(0, n, nq)
    q = np.zeros(nq, capi.PROJ_DTYPE)
    q["u"] = k["x"][src] + r.normal(0, 3, nq).astype(np.float32)
    q["v"] = k["y"][src] + r.normal(0, 3, nq).astype(np.float32)
    q["radius"] = r.choice([4.0, 7.0, 15.0, 30.0], nq).astype(np.float32) * scale[k["octave"][src]]
    lvl = k["octave"][src]
    q["min_level"] = np.maximum(lvl - 1, 0)
    q["max_level"] = lvl
    q["u_right"] = q["u"] - 20.0
    q["max_right_err"] = np.where(r.random(nq) < 0.5, 6.0, -1.0)
    q["angle"] = (k["angle"][src] + r.normal(0, 8, nq)) % 360
    q["level_lo"] = -1
    q["level_hi"] = -1
    qd = d[src].copy()
    flips = r.integers(0, 256, (nq, 12))
    for j in range(12):                                   # noisy copies of the source descriptors
        use = r.random(nq) < 0.6
        qd[use, flips[use, j] // 8] ^= (1 << (flips[use, j] % 8)).astype(np.uint8)
    if dup:
        qd[::3] = qd[0]
    u_right = np.where(r.random(n) < 0.6, k["x"] - 20.0 + r.normal(0, 4, n), -1.0).astype(np.float32)
    occupied = (r.random(n) < 0.1).astype(np.uint8)
    blocks = (r.random(nq) < 0.7).astype(np.uint8)
    return k, d, q, qd, u_right, occupied, blocks, scale


@pytest.mark.parametrize("variant,th,ratio,ori", [
    (capi.PROJ_LOCALMAP, 100, 0.8, False), (capi.PROJ_LASTFRAME, 100, 0.9, True),
    (capi.PROJ_KEYFRAME, 64, 0.9, True), (capi.PROJ_SIM3, 50, 0.75, False),
    (capi.PROJ_LASTFRAME, 100, 0.9, False), (capi.PROJ_LOCALMAP, 100, 0.99, True)])
def test_search_by_projection(variant, th, ratio, ori):
    for seed, dup in [(1, False), (2, True)]:
        k, d, q, qd, ur, occ, blk, sf = _frame_with_queries(seed, dup=dup)
        if variant in (capi.PROJ_KEYFRAME, capi.PROJ_SIM3):
            q = q.copy()
            q["min_level"] = -1
            q["max_level"] = -1
            q["level_lo"] = np.maximum(k["octave"][:len(q)] - 1, 0) if len(k) >= len(q) else 0
            q["level_hi"] = q["level_lo"] + 1
        fo = O.FrameData(k, d, 640, 480, u_right=ur, scale_factors=sf)
        fg = capi.FrameData(k, d, 640, 480, u_right=ur, scale_factors=sf)
        n_o, a_o, bi_o, bd_o, occ_o = O.search_by_projection(variant, fo, q, qd, blk, occ, th, ratio, ori)
        n_g, a_g, bi_g, bd_g, occ_g = capi.search_by_projection(variant, fg, q, qd, blk, occ, th, ratio, ori)
        assert n_g == n_o
        assert np.array_equal(bi_g, bi_o) and np.array_equal(bd_g, bd_o)
        assert np.array_equal(a_g, a_o) and np.array_equal(occ_g, occ_o)
        assert n_o > 50


def _random_frame(n, seed, w=640, h=480):
    """n keypoints at random places with random descriptors (nothing extracted: sizes the extractor never returns)"""
    r = np.random.default_rng(seed)
    k = np.zeros(n, capi.KP_DTYPE)
    k["x"] = r.uniform(0, w, n).astype(np.float32); k["y"] = r.uniform(0, h, n).astype(np.float32)
    k["octave"] = r.integers(0, 8, n); k["angle"] = r.uniform(0, 360, n).astype(np.float32); k["size"] = 31
    d = r.integers(0, 256, (n, 32), dtype=np.uint8)
    return k, d


def _queries_around(k, d, nq, seed, radius):
    r = np.random.default_rng(seed + 100)
    n = len(k)
    src = r.integers(0, n, nq)
    q = np.zeros(nq, capi.PROJ_DTYPE)
    q["u"] = k["x"][src] + r.normal(0, 2, nq).astype(np.float32)
    q["v"] = k["y"][src] + r.normal(0, 2, nq).astype(np.float32)
    q["radius"] = radius
    q["min_level"] = -1; q["max_level"] = -1; q["u_right"] = -1; q["max_right_err"] = -1
    q["angle"] = (k["angle"][src] + r.normal(0, 8, nq)) % 360
    q["level_lo"] = -1; q["level_hi"] = -1
    qd = d[src].copy()
    qd[:, 0] ^= r.integers(0, 256, nq, dtype=np.uint8)
    return q, qd, (r.random(n) < 0.1).astype(np.uint8), (r.random(nq) < 0.8).astype(np.uint8)


@pytest.mark.parametrize("n,nq,radius,variant,ratio,ori", [
    (1500, 6000, 70.0, capi.PROJ_LOCALMAP, 0.8, False),     # windows of > 32 keypoints that fill up: truncated lists are searched again
    (1500, 6000, 70.0, capi.PROJ_LASTFRAME, 0.9, True),
    (900, 3000, 1000.0, capi.PROJ_KEYFRAME, 0.9, True),     # every query sees the whole frame: long chains of conflicts
    (5000, 3000, 25.0, capi.PROJ_LOCALMAP, 0.8, False),     # more keypoints than the latency form takes: grid + warp replay
    (5000, 3000, 25.0, capi.PROJ_LASTFRAME, 0.9, True)])
def test_search_by_projection_contended(n, nq, radius, variant, ratio, ori):
    k, d = _random_frame(n, n + nq)
    q, qd, occ, blk = _queries_around(k, d, nq, n, radius)
    fo, fg = O.FrameData(k, d, 640, 480), capi.FrameData(k, d, 640, 480)
    n_o, a_o, bi_o, bd_o, occ_o = O.search_by_projection(variant, fo, q, qd, blk, occ, 100, ratio, ori)
    n_g, a_g, bi_g, bd_g, occ_g = capi.search_by_projection(variant, fg, q, qd, blk, occ, 100, ratio, ori)
    assert n_g == n_o and np.array_equal(bi_g, bi_o) and np.array_equal(bd_g, bd_o)
    assert np.array_equal(a_g, a_o) and np.array_equal(occ_g, occ_o)
    assert n_o > 100


def test_search_by_projection_empty():
    k = np.zeros(0, capi.KP_DTYPE)
    f = capi.FrameData(k, np.zeros((0, 32), np.uint8), 640, 480)
    q = np.zeros(3, capi.PROJ_DTYPE)
    q["radius"] = 5
    n, a, bi, bd, occ = capi.search_by_projection(0, f, q, np.zeros((3, 32), np.uint8), None, np.zeros(0, np.uint8), 100)
    assert n == 0 and bi.tolist() == [-1, -1, -1] and bd.tolist() == [256, 256, 256]


def _stereo_case(w, h, nf, seed):
    left, right = synth.stereo_pair(w, h, seed=seed)
    el, er = O.Extractor(nf), O.Extractor(nf)
    kl, dl = el.extract(left)
    kr, dr = er.extract(right)
    pl = [el.level(l) for l in range(8)]
    pr = [er.level(l) for l in range(8)]
    return left, right, kl, dl, kr, dr, pl, pr, el.tables()


@pytest.mark.parametrize("w,h,nf,seed", [(640, 240, 800, 2), (1241, 376, 2000, 5)])
def test_compute_stereo_matches_stateless(w, h, nf, seed):
    _, _, kl, dl, kr, dr, pl, pr, t = _stereo_case(w, h, nf, seed)
    bf, fx = 386.1448, 718.856
    u_o, z_o, c_o = O.compute_stereo_matches(kl, dl, kr, dr, pl, pr, t["scale"], t["inv_scale"], bf, fx)
    u_g, z_g, c_g = capi.compute_stereo_matches(kl, dl, kr, dr, pl, pr, t["scale"], t["inv_scale"], bf, fx)
    assert c_g == c_o
    assert np.array_equal(u_g, u_o) and np.array_equal(z_g, z_o)
    assert (u_o >= 0).sum() > 100
    # fewer left than right keypoints (found by the randomised soak: the kernel fills uRight / depth up to the larger count, which
    # must not spill into what follows them on the device) and the eyes the other way round
    for a, b in ((slice(0, len(kl) // 2), slice(None)), (slice(None), slice(0, len(kr) // 3))):
        for swap in (False, True):
            args = (kr[b], dr[b], kl[a], dl[a], pr, pl) if swap else (kl[a], dl[a], kr[b], dr[b], pl, pr)
            u_o, z_o, c_o = O.compute_stereo_matches(*args, t["scale"], t["inv_scale"], bf, fx)
            u_g, z_g, c_g = capi.compute_stereo_matches(*args, t["scale"], t["inv_scale"], bf, fx)
            assert c_g == c_o and np.array_equal(u_g, u_o) and np.array_equal(z_g, z_o)


def test_extract_stereo_batch_fused():
    w, h, nf, npairs = 1241, 376, 2000, 3
    bf, fx = 386.1448, 718.856
    L = np.stack([synth.stereo_pair(w, h, seed=10 + i)[0] for i in range(npairs)])
    R = np.stack([synth.stereo_pair(w, h, seed=10 + i)[1] for i in range(npairs)])
    ex = capi.Extractor(nfeatures=nf, max_width=w, max_height=h, chunk_frames=4, result_slots=2 * npairs)
    ur, dp, ncmp = ex.extract_stereo_batch(L, R, bf, fx)
    kps, desc, counts = ex.download(0, 2 * npairs)
    total = 0
    for i in range(npairs):
        el, er = O.Extractor(nf), O.Extractor(nf)
        kl, dl = el.extract(L[i])
        kr, dr = er.extract(R[i])
        assert counts[2 * i] == len(kl) and counts[2 * i + 1] == len(kr)
        assert kps[2 * i, :len(kl)].tobytes() == kl.tobytes() and kps[2 * i + 1, :len(kr)].tobytes() == kr.tobytes()
        assert np.array_equal(desc[2 * i, :len(kl)], dl) and np.array_equal(desc[2 * i + 1, :len(kr)], dr)
        t = el.tables()
        u_o, z_o, c_o = O.compute_stereo_matches(kl, dl, kr, dr, [el.level(l) for l in range(8)],
                                                 [er.level(l) for l in range(8)], t["scale"], t["inv_scale"], bf, fx)
        total += c_o
        assert np.array_equal(ur[i, :len(kl)], u_o) and np.array_equal(dp[i, :len(kl)], z_o), i
    assert ncmp == total
    # device-resident images (the bench path) give the same answer
    Ld, Rd = torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda()
    ur2, dp2, _ = ex.extract_stereo_batch(Ld.data_ptr(), Rd.data_ptr(), bf, fx, nframes=npairs, width=w, height=h,
                                          frame_stride=w * h, row_stride=w, on_device=True)
    assert np.array_equal(ur2, ur) and np.array_equal(dp2, dp)
    ex.close()


def _knn_gpu(q, db, offset=0, variant=1):
    capi.lib().orbm_knn2_set_variant(variant)
    qd, dd = torch.from_numpy(q).cuda(), torch.from_numpy(db).cuda()
    od = torch.empty((len(q), 2), dtype=torch.int16, device="cuda")
    oi = torch.empty((len(q), 2), dtype=torch.int32, device="cuda")
    capi.knn2_device(qd.data_ptr(), len(q), dd.data_ptr(), len(db), offset, od.data_ptr(), oi.data_ptr(),
                     torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return od.cpu().numpy().view(np.uint16), oi.cpu().numpy()


@pytest.mark.parametrize("nq,ndb,variant", [(97, 1500, 0), (97, 1500, 1), (97, 1500, 2), (4096, 65536, 1), (1, 1, 1),
                                            (1025, 300, 2), (5000, 9000, 0)])
def test_knn2_matches_oracle(nq, ndb, variant):
    q = synth.random_descriptors(nq, 1)
    db = synth.random_descriptors(ndb, 2)
    if ndb > 800:
        db[700] = db[20]
        q = synth.plant_near_duplicates(q, db, frac=0.05)
        q[5 % nq] = db[20]
    od_o, oi_o = O.knn2(q, db, offset=1000, nthreads=8)
    od_g, oi_g = _knn_gpu(q, db, offset=1000, variant=variant)
    assert np.array_equal(oi_g, oi_o)
    assert np.array_equal(od_g, od_o)


def test_knn2_sharded_merge_equals_whole():
    nq, ndb, shards = 3000, 40000, 8
    q = synth.plant_near_duplicates(synth.random_descriptors(nq, 3), synth.random_descriptors(ndb, 4), frac=0.02)
    db = synth.random_descriptors(ndb, 4)
    od_w, oi_w = _knn_gpu(q, db)
    per = ndb // shards
    pd = torch.empty((shards, nq, 2), dtype=torch.int16, device="cuda")
    pi = torch.empty((shards, nq, 2), dtype=torch.int32, device="cuda")
    qd = torch.from_numpy(q).cuda()
    keep = []
    for s in range(shards):
        dd = torch.from_numpy(db[s * per:(s + 1) * per]).cuda()
        keep.append(dd)
        capi.knn2_device(qd.data_ptr(), nq, dd.data_ptr(), per, s * per, pd[s].data_ptr(), pi[s].data_ptr(),
                         torch.cuda.current_stream().cuda_stream)
    od = torch.empty((nq, 2), dtype=torch.int16, device="cuda")
    oi = torch.empty((nq, 2), dtype=torch.int32, device="cuda")
    capi.knn2_merge_device(pd.data_ptr(), pi.data_ptr(), shards, nq, od.data_ptr(), oi.data_ptr(),
                           torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert np.array_equal(oi.cpu().numpy(), oi_w) and np.array_equal(od.cpu().numpy().view(np.uint16), od_w)
    od_o, oi_o = O.knn2(q[:200], db, nthreads=8)
    assert np.array_equal(oi_w[:200], oi_o) and np.array_equal(od_w[:200], od_o)


def test_stateless_calls_reuse_scratch_across_sizes_and_threads():
    """The stateless matcher entry points draw device scratch from a per-thread block cache (orbx_scratch_alloc): the same
    calls repeated with growing and shrinking inputs, interleaved between two threads, must keep returning the oracle's
    answers (stale contents of a reused block, or a block shared between threads, would show up here)."""
    import threading
    cases = []
    for seed, nq in ((1, 900), (2, 300), (3, 1200), (4, 60)):
        k, d, q, qd, ur, occ, blk, sf = _frame_with_queries(seed, nq=nq)
        fo = O.FrameData(k, d, 640, 480, u_right=ur, scale_factors=sf)
        want = O.search_by_projection(capi.PROJ_LOCALMAP, fo, q, qd, blk, occ, 100, 0.8, False)
        cases.append((k, d, q, qd, ur, occ, blk, sf, want))
    errors = []

    def worker(order):
        try:
            for rep in range(3):
                for ci in order:
                    k, d, q, qd, ur, occ, blk, sf, want = cases[ci]
                    fg = capi.FrameData(k, d, 640, 480, u_right=ur, scale_factors=sf)
                    got = capi.search_by_projection(capi.PROJ_LOCALMAP, fg, q, qd, blk, occ, 100, 0.8, False)
                    assert got[0] == want[0]
                    for a, b in zip(got[1:], want[1:]):
                        assert np.array_equal(a, b)
                    dist = capi.descriptor_distance(d[: 50 + 10 * ci], d[::-1][: 50 + 10 * ci])
                    ref = np.unpackbits(d[: 50 + 10 * ci] ^ d[::-1][: 50 + 10 * ci], axis=1).sum(axis=1)
                    assert np.array_equal(dist, ref)
        except Exception as e:      # surfaced in the main thread
            errors.append(e)

    ts = [threading.Thread(target=worker, args=(o,)) for o in ([0, 1, 2, 3], [3, 2, 1, 0])]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert not errors, errors[0]


def test_extract_stereo_batch_padded_host_rows():
    """contiguous host frames whose row stride exceeds the workspace pitch (1280 for width 1241): each eye
    owns half of the landing zone, the linear copy of one eye must not spill into the other."""
    w, h, nf, npairs, stride = 1241, 376, 2000, 2, 1280
    bf, fx = 386.1448, 718.856
    Lp = np.zeros((npairs, h, stride), np.uint8); Rp = np.zeros((npairs, h, stride), np.uint8)
    for i in range(npairs):
        l, r = synth.stereo_pair(w, h, seed=40 + i)
        Lp[i, :, :w] = l; Rp[i, :, :w] = r
        Lp[i, :, w:] = 255; Rp[i, :, w:] = 7
    ex = capi.Extractor(nfeatures=nf, max_width=w, max_height=h, chunk_frames=4, result_slots=2 * npairs)
    ur, dp, _ = ex.extract_stereo_batch(Lp.ctypes.data, Rp.ctypes.data, bf, fx, nframes=npairs, width=w, height=h,
                                        frame_stride=h * stride, row_stride=stride, on_device=False)
    kps, desc, counts = ex.download(0, 2 * npairs)
    for i in range(npairs):
        el, er = O.Extractor(nf), O.Extractor(nf)
        kl, dl = el.extract(np.ascontiguousarray(Lp[i, :, :w]))
        kr, dr = er.extract(np.ascontiguousarray(Rp[i, :, :w]))
        assert counts[2 * i] == len(kl) and counts[2 * i + 1] == len(kr)
        assert kps[2 * i, :len(kl)].tobytes() == kl.tobytes() and kps[2 * i + 1, :len(kr)].tobytes() == kr.tobytes()
        assert np.array_equal(desc[2 * i, :len(kl)], dl) and np.array_equal(desc[2 * i + 1, :len(kr)], dr)
        t = el.tables()
        u_o, z_o, _ = O.compute_stereo_matches(kl, dl, kr, dr, [el.level(l) for l in range(8)],
                                               [er.level(l) for l in range(8)], t["scale"], t["inv_scale"], bf, fx)
        assert np.array_equal(ur[i, :len(kl)], u_o) and np.array_equal(dp[i, :len(kl)], z_o), i
    ex.close()
