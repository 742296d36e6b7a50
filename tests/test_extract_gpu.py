"""GPU parity: liborbx.so (through the C ABI) vs the CPU oracle, stage by stage and end to end.
Bit-exact for pyramid bytes, FAST candidates, culled keypoints (x, y, octave, order), descriptors;
angle/response are compared exactly too (the tolerance north_star allows is 1e-5 relative)."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as O
from orb_slam2_map_b200 import capi, synth

pytestmark = pytest.mark.gpu

SHAPES = [(640, 480, 1000), (1241, 376, 2000), (752, 480, 1200), (333, 251, 500), (1920, 1080, 5000)]


def _image(w, h, seed):
    return synth.textured_canvas(w, h, seed=seed)


def _compare_full(kg, dg, ko, do):
    assert len(kg) == len(ko), (len(kg), len(ko))
    for f in ("x", "y", "size", "octave", "class_id", "response"):
        assert np.array_equal(kg[f], ko[f]), f
    assert np.array_equal(kg["angle"], ko["angle"]), np.abs(kg["angle"] - ko["angle"]).max()
    assert np.array_equal(dg, do), int((dg != do).any(axis=1).sum())


@pytest.mark.parametrize("w,h,nf", SHAPES)
def test_stages_and_full(w, h, nf):
    img = _image(w, h, seed=w + h)
    ex = capi.Extractor(nfeatures=nf, max_width=w, max_height=h)
    orc = O.Extractor(nfeatures=nf)
    ko, do = orc.extract(img)
    kg, dg = ex.extract(img)
    # tables
    tg, to = ex.tables(), orc.tables()
    for k in tg:
        assert np.array_equal(tg[k], to[k]), k
    # pyramid + blur bytes
    for l in range(8):
        assert np.array_equal(ex.level(0, l), orc.level(l)), f"raw level {l}"
        assert np.array_equal(ex.level(0, l, blurred=True), orc.level(l, blurred=True)), f"blur level {l}"
    assert np.array_equal(ex.level(0, 2, border=19), O.border101(orc.level(2), 19))
    # FAST candidates in reference order
    for l in range(8):
        cg, co = ex.candidates(0, l), orc.candidates(l)
        assert cg.shape == co.shape, (l, cg.shape, co.shape)
        assert np.array_equal(cg, co), f"candidates level {l}"
    _compare_full(kg, dg, ko, do)
    ex.close()


@pytest.mark.parametrize("kind", ["noise", "flat", "checker", "gradient", "sparse"])
def test_odd_content(kind):
    w, h = 640, 480
    rng = np.random.default_rng(5)
    if kind == "noise":
        img = rng.integers(0, 256, (h, w), dtype=np.uint8)
    elif kind == "flat":
        img = np.full((h, w), 77, np.uint8)
    elif kind == "checker":
        yy, xx = np.mgrid[0:h, 0:w]
        img = (((yy // 9 + xx // 7) % 2) * 200 + 20).astype(np.uint8)
    elif kind == "gradient":
        yy, xx = np.mgrid[0:h, 0:w]
        img = ((xx * 255) // w).astype(np.uint8)
    else:
        img = np.full((h, w), 128, np.uint8)
        pts = rng.integers(30, 450, (40, 2))
        for (y, x) in pts:
            img[y:y + 5, x:x + 5] = 255
    ex = capi.Extractor(max_width=w, max_height=h)
    orc = O.Extractor()
    ko, do = orc.extract(img)
    kg, dg = ex.extract(img)
    for l in range(8):
        assert np.array_equal(ex.candidates(0, l), orc.candidates(l)), f"candidates level {l}"
    _compare_full(kg, dg, ko, do)
    ex.close()


def test_batch_matches_single_and_oracle():
    w, h, n = 640, 480, 12
    frames = synth.frame_sequence(n, w, h, seed=3)
    ex = capi.Extractor(max_width=w, max_height=h, chunk_frames=5, result_slots=n)
    ex.extract_batch(frames)
    kps, desc, counts = ex.download(0, n)
    orc = O.Extractor()
    for i in range(n):
        ko, do = orc.extract(frames[i])
        _compare_full(kps[i, :counts[i]], desc[i, :counts[i]], ko, do)
    ex.close()


def test_extract_batch_to_host_overlapped_copies():
    """host in / host out with chunked, overlapped copies == device-resident results == oracle"""
    w, h, n = 640, 480, 11
    frames = synth.frame_sequence(n, w, h, seed=8)
    ex = capi.Extractor(max_width=w, max_height=h, chunk_frames=3, result_slots=n)
    kps = np.zeros((n, ex.cap), capi.KP_DTYPE)
    desc = np.zeros((n, ex.cap, 32), np.uint8)
    counts = np.zeros(n, np.int32)
    for _ in range(2):                                   # second call reuses both staging buffers
        ex.extract_batch_to_host(frames, kps, desc, counts)
    k2, d2, c2 = ex.download(0, n)
    assert np.array_equal(counts, c2)
    orc = O.Extractor()
    for i in range(n):
        ko, do = orc.extract(frames[i])
        _compare_full(kps[i, :counts[i]], desc[i, :counts[i]], ko, do)
        assert kps[i, :counts[i]].tobytes() == k2[i, :c2[i]].tobytes() and np.array_equal(desc[i, :counts[i]], d2[i, :c2[i]])
    ex.close()


def test_params_variants():
    img = _image(640, 480, seed=11)
    for (nf, sc, nl, it, mt) in [(500, 1.2, 8, 20, 7), (1500, 1.3, 6, 30, 10), (300, 1.5, 4, 15, 5),
                                 (2000, 1.1, 10, 20, 7)]:
        ex = capi.Extractor(nf, sc, nl, it, mt, max_width=640, max_height=480)
        orc = O.Extractor(nf, sc, nl, it, mt)
        ko, do = orc.extract(img)
        kg, dg = ex.extract(img)
        _compare_full(kg, dg, ko, do)
        ex.close()


def test_threshold_order_variants():
    """cv::FAST(iniThFAST) first, cv::FAST(minThFAST) only for empty cells (ORBextractor.cc:809-816) — also when
    ini <= min, and above the 7-bit range of the byte-packed pre-test"""
    img = _image(640, 480, seed=21)
    for (it, mt) in [(5, 7), (7, 7), (60, 3), (126, 100), (140, 130)]:
        ex = capi.Extractor(1000, 1.2, 8, it, mt, max_width=640, max_height=480)
        orc = O.Extractor(1000, 1.2, 8, it, mt)
        ko, do = orc.extract(img)
        kg, dg = ex.extract(img)
        for l in range(8):
            assert np.array_equal(ex.candidates(0, l), orc.candidates(l)), f"candidates level {l} ini={it} min={mt}"
        _compare_full(kg, dg, ko, do)
        ex.close()


def test_octree_global_key_path(monkeypatch):
    """levels with more candidates than the shared-memory key capacity run out of the global key buffers"""
    monkeypatch.setenv("ORBX_OCTREE_KEYCAP", "64")
    rng = np.random.default_rng(9)
    for img in (_image(640, 480, seed=31), rng.integers(0, 256, (480, 640), dtype=np.uint8)):
        ex = capi.Extractor(max_width=640, max_height=480)
        orc = O.Extractor()
        ko, do = orc.extract(img)
        kg, dg = ex.extract(img)
        _compare_full(kg, dg, ko, do)
        ex.close()


def test_tall_and_wide_images():
    """several quadtree roots (wide) / a single root with tall cells (tall), small levels vanish"""
    for (w, h, nf) in [(1200, 160, 800), (160, 900, 600), (97, 83, 200)]:
        img = _image(w, h, seed=w * 3 + h)
        ex = capi.Extractor(nfeatures=nf, max_width=w, max_height=h)
        orc = O.Extractor(nfeatures=nf)
        ko, do = orc.extract(img)
        kg, dg = ex.extract(img)
        for l in range(8):
            assert np.array_equal(ex.candidates(0, l), orc.candidates(l)), f"candidates level {l} {w}x{h}"
        _compare_full(kg, dg, ko, do)
        ex.close()


def test_empty_image_is_silent():
    ex = capi.Extractor()
    n = C.c_int(-7)
    rc = capi.lib().orbx_extract(ex.h, None, 0, 0, 0, None, None, 0, C.byref(n))
    assert rc == 0 and n.value == -7
    ex.close()


def test_single_frame_graph_replay(monkeypatch):
    """orbx_extract on ONE handle, call after call (the way Tracking uses ORBextractor::operator()): the first call of a
    shape captures the frame's kernels into a CUDA graph, later calls replay it.  Every call must equal the oracle bit
    for bit: different content per call, a strided source, a shape change (the graph is dropped with the workspace) and
    back, batch calls in between, pyramid levels after a replay, and the plain-launch path (ORBX_NO_GRAPH)."""
    orc = O.Extractor()
    ex = capi.Extractor(max_width=752, max_height=480, chunk_frames=4, result_slots=4)
    seq = synth.frame_sequence(6, 640, 480, seed=11)
    for i in range(6):                               # call 0 plain + capture, 1.. replays
        kg, dg = ex.extract(seq[i])
        ko, do = orc.extract(seq[i])
        _compare_full(kg, dg, ko, do)
    for l in (0, 3, 7):                              # mvImagePyramid of the frame the graph just processed
        assert np.array_equal(ex.level(0, l), orc.level(l))
        assert np.array_equal(ex.level(0, l, blurred=True), orc.level(l, blurred=True))
    wide = np.zeros((480, 700), np.uint8)            # row stride != width
    wide[:, :640] = seq[2]
    wide[:, 640:] = 255
    kg, dg = ex.extract(wide[:, :640])
    ko, do = orc.extract(seq[2])
    _compare_full(kg, dg, ko, do)
    other = _image(752, 480, seed=3)                 # another shape: workspace and graph rebuilt
    for _ in range(2):
        kg, dg = ex.extract(other)
        ko, do = orc.extract(other)
        _compare_full(kg, dg, ko, do)
    ex.extract_batch(seq[:4])                        # a batch on the same handle, then single frames again
    kb, db, cb = ex.download(0, 4)
    for i in range(4):
        ko, do = orc.extract(seq[i])
        _compare_full(kb[i][:cb[i]], db[i][:cb[i]], ko, do)
    for i in (5, 1, 4):
        kg, dg = ex.extract(seq[i])
        ko, do = orc.extract(seq[i])
        _compare_full(kg, dg, ko, do)
    import torch
    pin = torch.empty((480, 640), dtype=torch.uint8, pin_memory=True)      # page-locked source: copied in place, no staging
    pin.copy_(torch.from_numpy(seq[3]))
    kg, dg = ex.extract(pin.numpy())
    ko, do = orc.extract(seq[3])
    _compare_full(kg, dg, ko, do)
    flat = np.full((480, 640), 9, np.uint8)          # no keypoints at all after a full frame
    kg, dg = ex.extract(flat)
    assert len(kg) == 0 and dg.shape == (0, 32)
    kg, dg = ex.extract(seq[0])
    ko, do = orc.extract(seq[0])
    _compare_full(kg, dg, ko, do)
    monkeypatch.setenv("ORBX_NO_GRAPH", "1")
    kg2, dg2 = ex.extract(seq[0])
    _compare_full(kg2, dg2, ko, do)
    ex.close()


def test_wide_image_on_larger_handle_grows_capacity():
    """a smaller image with a wider aspect ratio than the handle maximum has more quadtree roots per level
    (nIni, src/ORBextractor.cc:543-546) and so a larger keypoint capacity than the one sized at creation: the result
    store grows (orbx_reserve_shape reports it) instead of the call failing with ORBX_E_CAPACITY."""
    ex = capi.Extractor(nfeatures=100, max_width=1920, max_height=1080, chunk_frames=2, result_slots=2)
    cap0 = ex.cap
    for (w, h) in [(1241, 100), (640, 480), (1900, 70)]:
        img = _image(w, h, seed=w + 7 * h)
        orc = O.Extractor(nfeatures=100)
        ko, do = orc.extract(img)
        kg, dg = ex.extract(img)
        _compare_full(kg, dg, ko, do)
        ex.extract_batch(np.stack([img, img]))           # the batch path on the grown store
        kb, db, cb = ex.download(0, 2)
        for i in range(2):
            _compare_full(kb[i][:cb[i]], db[i][:cb[i]], ko, do)
    assert ex.cap > cap0
    ex.close()


def test_failed_shape_does_not_poison_the_handle():
    """a shape the handle rejects must leave no trace: the next call with the previous shape runs on a
    workspace laid out for THAT shape (the handle forgets its shape before it touches a buffer)."""
    ex = capi.Extractor(max_width=640, max_height=480)
    orc = O.Extractor()
    img = _image(640, 480, seed=21)
    ko, do = orc.extract(img)
    _compare_full(*ex.extract(img), ko, do)
    with pytest.raises(capi.OrbxError):
        ex.extract(_image(800, 480, seed=1))             # exceeds the handle maximum
    _compare_full(*ex.extract(img), ko, do)
    small = _image(333, 251, seed=2)
    ks, ds = O.Extractor().extract(small)
    _compare_full(*ex.extract(small), ks, ds)
    _compare_full(*ex.extract(img), ko, do)
    ex.close()
