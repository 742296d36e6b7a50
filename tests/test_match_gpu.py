"""GPU parity of the Frame grid + ORBmatcher searches (through the C ABI) against the CPU oracle.
Match indices, counts and distances must be identical."""
import numpy as np
import pytest

import oracle_lib as O
from orb_slam2_map_b200 import capi, synth

pytestmark = pytest.mark.gpu


def _frames(n, w=640, h=480, seed=3):
    return synth.frame_sequence(n, w, h, seed=seed)


def test_descriptor_distance_exhaustive_kats():
    rng = np.random.default_rng(0)
    a = rng.integers(0, 256, (4096, 32), dtype=np.uint8)
    b = rng.integers(0, 256, (4096, 32), dtype=np.uint8)
    a[0] = 0; b[0] = 0            # 0
    a[1] = 0; b[1] = 255          # 256
    a[2] = 0; b[2] = 0; b[2, 17] = 0x10   # 1
    got = capi.descriptor_distance(a, b)
    exp = np.array([O.descriptor_distance(a[i], b[i]) for i in range(len(a))], np.int32)
    assert np.array_equal(got, exp)
    assert got[0] == 0 and got[1] == 256 and got[2] == 1
    assert np.array_equal(got, np.unpackbits(a ^ b, axis=1).sum(axis=1))


def test_search_init_pairs_on_device_slots():
    w, h, n = 640, 480, 9
    frames = _frames(n)
    ex = capi.Extractor(max_width=w, max_height=h, chunk_frames=4, result_slots=n)
    ex.extract_batch(frames)
    kps, desc, counts = ex.download(0, n)
    m12, nm, ncmp = ex.search_init_pairs(0, n - 1, window=100, nnratio=0.9, check_ori=True)
    total_cmp = 0
    for i in range(n - 1):
        f1 = O.FrameData(kps[i, :counts[i]], desc[i, :counts[i]], w, h)
        f2 = O.FrameData(kps[i + 1, :counts[i + 1]], desc[i + 1, :counts[i + 1]], w, h)
        prev = np.stack([f1.kps["x"], f1.kps["y"]], axis=1)
        n_o, m_o, _, c_o = O.search_for_initialization(f1, f2, prev, 100, 0.9, True)
        total_cmp += c_o
        assert nm[i] == n_o, (i, nm[i], n_o)
        assert np.array_equal(m12[i, :counts[i]], m_o), i
        assert n_o > 20      # the synthetic sequence really matches
    assert ncmp == total_cmp
    ex.close()


@pytest.mark.parametrize("window,nnratio,ori", [(100, 0.9, True), (30, 0.7, False), (250, 0.95, True), (10, 0.9, True)])
def test_search_for_initialization_stateless(window, nnratio, ori):
    w, h = 640, 480
    frames = _frames(3, seed=7)
    orc = O.Extractor()
    k1, d1 = orc.extract(frames[0])
    k2, d2 = orc.extract(frames[2])
    prev = np.stack([k1["x"], k1["y"]], axis=1).astype(np.float32)
    # second round with updated prev_matched exercises the in/out argument
    fo1, fo2 = O.FrameData(k1, d1, w, h), O.FrameData(k2, d2, w, h)
    fg1, fg2 = capi.FrameData(k1, d1, w, h), capi.FrameData(k2, d2, w, h)
    n_o, m_o, p_o, _ = O.search_for_initialization(fo1, fo2, prev, window, nnratio, ori)
    n_g, m_g, p_g = capi.search_for_initialization(fg1, fg2, prev, window, nnratio, ori)
    assert n_g == n_o and np.array_equal(m_g, m_o) and np.array_equal(p_g, p_o)
    n_o2, m_o2, p_o2, _ = O.search_for_initialization(fo1, fo2, p_o, window, nnratio, ori)
    n_g2, m_g2, p_g2 = capi.search_for_initialization(fg1, fg2, p_g, window, nnratio, ori)
    assert n_g2 == n_o2 and np.array_equal(m_g2, m_o2) and np.array_equal(p_g2, p_o2)


def test_search_init_collisions_and_duplicates():
    """Many identical descriptors -> ties, steals and the vMatchedDistance skip all fire."""
    w, h = 640, 480
    rng = np.random.default_rng(1)
    n = 400
    base = rng.integers(0, 256, (8, 32), dtype=np.uint8)
    def mk(seed):
        r = np.random.default_rng(seed)
        k = np.zeros(n, capi.KP_DTYPE)
        k["x"] = r.integers(20, 620, n); k["y"] = r.integers(20, 460, n)
        k["octave"] = r.integers(0, 2, n); k["angle"] = r.uniform(0, 360, n).astype(np.float32)
        k["size"] = 31; k["class_id"] = -1
        d = base[r.integers(0, 8, n)].copy()
        flip = r.integers(0, 32, n)
        d[np.arange(n), flip] ^= (1 << r.integers(0, 8, n)).astype(np.uint8)
        return k, d
    k1, d1 = mk(2)
    k2, d2 = mk(3)
    prev = np.stack([k1["x"], k1["y"]], axis=1).astype(np.float32)
    for (bounds) in [None, (-10.5, 660.25, -7.0, 500.0)]:
        fo1, fo2 = O.FrameData(k1, d1, w, h, bounds=bounds), O.FrameData(k2, d2, w, h, bounds=bounds)
        fg1, fg2 = capi.FrameData(k1, d1, w, h, bounds=bounds), capi.FrameData(k2, d2, w, h, bounds=bounds)
        n_o, m_o, p_o, _ = O.search_for_initialization(fo1, fo2, prev, 120, 0.99, True)
        n_g, m_g, p_g = capi.search_for_initialization(fg1, fg2, prev, 120, 0.99, True)
        assert n_g == n_o and np.array_equal(m_g, m_o) and np.array_equal(p_g, p_o)


@pytest.mark.parametrize("n,window,nnratio", [(3000, 300, 0.99), (1200, 640, 0.9), (5000, 120, 0.99)])
def test_search_init_contended_and_large_frames(n, window, nnratio):
    """Thousands of near-identical descriptors inside wide windows: the 32-entry candidate lists of the latency form fill up
    with ties and are searched again as keypoints are taken; 5000 keypoints take the batch kernel (grid + warp replay)."""
    w, h = 640, 480
    base = np.random.default_rng(n).integers(0, 256, (6, 32), dtype=np.uint8)
    def mk(seed):
        r = np.random.default_rng(seed)
        k = np.zeros(n, capi.KP_DTYPE)
        k["x"] = r.uniform(5, 635, n).astype(np.float32); k["y"] = r.uniform(5, 475, n).astype(np.float32)
        k["octave"] = r.integers(0, 2, n); k["angle"] = r.uniform(0, 360, n).astype(np.float32)
        k["size"] = 31; k["class_id"] = -1
        d = base[r.integers(0, 6, n)].copy()
        for _ in range(3):
            d[np.arange(n), r.integers(0, 32, n)] ^= (1 << r.integers(0, 8, n)).astype(np.uint8)
        return k, d
    k1, d1 = mk(n + 1)
    k2, d2 = mk(n + 2)
    prev = np.stack([k1["x"], k1["y"]], axis=1).astype(np.float32)
    fo1, fo2 = O.FrameData(k1, d1, w, h), O.FrameData(k2, d2, w, h)
    fg1, fg2 = capi.FrameData(k1, d1, w, h), capi.FrameData(k2, d2, w, h)
    for ori in (True, False):
        n_o, m_o, p_o, _ = O.search_for_initialization(fo1, fo2, prev, window, nnratio, ori)
        n_g, m_g, p_g = capi.search_for_initialization(fg1, fg2, prev, window, nnratio, ori)
        assert n_g == n_o and np.array_equal(m_g, m_o) and np.array_equal(p_g, p_o)
    assert n_o > 20


def test_search_init_empty_frames():
    k = np.zeros(0, capi.KP_DTYPE)
    d = np.zeros((0, 32), np.uint8)
    f = capi.FrameData(k, d, 640, 480)
    n_g, m_g, _ = capi.search_for_initialization(f, f, np.zeros((0, 2), np.float32))
    assert n_g == 0 and len(m_g) == 0


def test_extract_match_pipeline_equals_separate_calls():
    """orbx_extract_match_to_host (matcher folded into the chunk pipeline, ramped and tapered chunks) against
    batch extraction followed by orbx_search_init_pairs."""
    n = 23
    frames = synth.frame_sequence(n, 640, 480, seed=5)
    ex = capi.Extractor(max_width=640, max_height=480, chunk_frames=4, result_slots=n)
    cap = ex.cap
    kp = np.zeros((n, cap), capi.KP_DTYPE); de = np.zeros((n, cap, 32), np.uint8); cnt = np.zeros(n, np.int32)
    m12 = np.full((n - 1, cap), -7, np.int32); nm = np.full(n - 1, -7, np.int32)
    ex.extract_match_to_host(frames, kp, de, cnt, m12, nm, 100, 0.9, True)
    ex2 = capi.Extractor(max_width=640, max_height=480, chunk_frames=8, result_slots=n)
    ex2.extract_batch(frames)
    k2, d2, c2 = ex2.download(0, n)
    m2, nm2, _ = ex2.search_init_pairs(0, n - 1, 100, 0.9, True)
    assert np.array_equal(cnt, c2) and nm.min() > 50
    for f in range(n):
        assert kp[f, :cnt[f]].tobytes() == k2[f, :c2[f]].tobytes() and np.array_equal(de[f, :cnt[f]], d2[f, :c2[f]])
    assert np.array_equal(nm, nm2)
    for f in range(n - 1):
        assert np.array_equal(m12[f, :cnt[f]], m2[f, :c2[f]]), f"pair {f}"


def test_batch_pipeline_pageable_pool_equals_page_locked_buffers():
    """Pageable caller buffers go through the handle's page-locked pools (inputs copied in by host threads, results handed over a
    chunk later); page-locked buffers are copied directly.  Same frames, padded rows on the pageable side: identical results."""
    import torch
    n, w, h = 37, 640, 480
    frames = synth.frame_sequence(n, w, h, seed=9)
    ex = capi.Extractor(max_width=w, max_height=h, chunk_frames=8, result_slots=n)
    cap = ex.cap
    def run(pinned):
        if pinned:
            pin = lambda shape, dt: torch.empty(shape, dtype=dt, pin_memory=True).numpy()
            src = pin((n, h, w), torch.uint8); src[:] = frames
            kp = pin((n, cap, 28), torch.uint8).view(capi.KP_DTYPE).reshape(n, cap); de = pin((n, cap, 32), torch.uint8)
            cnt = pin((n,), torch.int32); m12 = pin((n - 1, cap), torch.int32); nm = pin((n - 1,), torch.int32)
        else:
            padded = np.zeros((n, h, w + 24), np.uint8); padded[:, :, :w] = frames
            src = padded[:, :, :w]                                   # row stride 664: the pool copies row by row
            kp = np.zeros((n, cap), capi.KP_DTYPE); de = np.zeros((n, cap, 32), np.uint8)
            cnt = np.zeros(n, np.int32); m12 = np.full((n - 1, cap), -7, np.int32); nm = np.full(n - 1, -7, np.int32)
        ex.extract_match_to_host(src, kp, de, cnt, m12, nm, 100, 0.9, True)
        return kp, de, cnt, m12, nm
    a, b = run(True), run(False)
    assert np.array_equal(a[2], b[2]) and np.array_equal(a[4], b[4]) and a[4].min() > 50
    for f in range(n):
        c = a[2][f]
        assert a[0][f, :c].tobytes() == b[0][f, :c].tobytes() and np.array_equal(a[1][f, :c], b[1][f, :c])
        if f < n - 1:
            assert np.array_equal(a[3][f, :c], b[3][f, :c])
