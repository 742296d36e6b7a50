"""K19-K21 (orbx_kfmatch.cu) through the C ABI against oracle/kfmatch_oracle.cpp: bit-exact indices, distances and counts."""
import numpy as np
import pytest

import oracle_lib as O
from orb_slam2_map_b200 import capi
from test_oracle_kfmatch import SCALE, SIGMA2, make_fuse_case, make_sim3_case, make_tri_pair

pytestmark = pytest.mark.gpu


def _to_capi(kf):
    """the same arrays behind a capi.KeyFrameData"""
    fd = kf.fd
    return capi.KeyFrameData(fd.kps, fd.desc, 640, 480, fd.u_right, fd.sf, kf.s2, kf.node, kf.flags, bounds=fd.bounds)


@pytest.mark.parametrize("variant", [0, 1, 2])
def test_fuse_variants(variant):
    total = 0
    for seed in (3, 4):
        kf, q, qd = make_fuse_case(seed * 7 + variant, n=2000, nq=3000)
        bo, do = O.fuse(variant, kf, q, qd)
        bg, dg = capi.fuse(variant, _to_capi(kf), q, qd)
        assert np.array_equal(bg, bo) and np.array_equal(dg, do)
        total += int((bg >= 0).sum())
    assert total > 500


def test_fuse_and_sim3_large_keyframes_take_the_grid_path():
    """more than 4096 keypoints: K20 walks the grid (the latency form sweeps the keypoints instead)"""
    kf, q, qd = make_fuse_case(5, n=5000, nq=2500)
    for variant in (0, 1):
        bo, do = O.fuse(variant, kf, q, qd)
        bg, dg = capi.fuse(variant, _to_capi(kf), q, qd)
        assert np.array_equal(bg, bo) and np.array_equal(dg, do)
    assert (bg >= 0).sum() > 200


def test_fuse_empty_and_mono_keyframe():
    kf, q, qd = make_fuse_case(77, n=300, nq=64)
    fd = kf.fd
    mono = O.KeyFrameData(fd.kps, fd.desc, 640, 480, None, SCALE, SIGMA2)
    bo, do = O.fuse(0, mono, q, qd)
    bg, dg = capi.fuse(0, _to_capi(mono), q, qd)
    assert np.array_equal(bg, bo) and np.array_equal(dg, do)
    empty = capi.KeyFrameData(fd.kps[:0], fd.desc[:0], 640, 480, None, SCALE, SIGMA2)
    bg, dg = capi.fuse(1, empty, q, qd)
    assert (bg == -1).all() and (dg == 256).all()
    bg, _ = capi.fuse(1, _to_capi(kf), q[:0], qd[:0])
    assert len(bg) == 0


def test_search_by_sim3():
    for seed in (21, 22):
        kf1, kf2, q12, d12, i1, q21, d21, i2 = make_sim3_case(seed)
        nf_o, m_o = O.search_by_sim3(kf1, kf2, q12, d12, i1, q21, d21, i2)
        nf_g, m_g = capi.search_by_sim3(_to_capi(kf1), _to_capi(kf2), q12, d12, i1, q21, d21, i2)
        assert nf_g == nf_o and np.array_equal(m_g, m_o)
        assert nf_o > 20


@pytest.mark.parametrize("only_stereo,ori", [(False, True), (True, True), (False, False)])
def test_search_for_triangulation_batch(only_stereo, ori):
    cases = [make_tri_pair(s, n1=n1, n2=n2, mono_frac=m) for s, n1, n2, m in
             [(1, 400, 520, 0.5), (2, 900, 1100, 0.3), (3, 33, 700, 1.0), (4, 2000, 1900, 0.0), (5, 1, 1, 0.5)]]
    pairs = [(_to_capi(a), _to_capi(b)) for a, b, _, _ in cases]
    nm, ms = capi.search_for_triangulation(pairs, [c[2] for c in cases], [c[3] for c in cases], only_stereo, ori)
    total = 0
    for p, (a, b, F, epi) in enumerate(cases):
        n_o, m_o = O.search_for_triangulation(a, b, F, epi, only_stereo, ori)
        assert nm[p] == n_o and np.array_equal(ms[p], m_o), f"pair {p}"
        total += n_o
    assert total > 100


def test_search_for_triangulation_large_keyframes_take_the_sort_path():
    """more than 4096 features in KF2: K19 sorts KF2's (node, index) keys instead of sweeping them"""
    a, b, F, epi = make_tri_pair(6, n1=4400, n2=5000, mono_frac=0.4)
    nm, ms = capi.search_for_triangulation([(_to_capi(a), _to_capi(b))], [F], [epi], False, True)
    n_o, m_o = O.search_for_triangulation(a, b, F, epi, False, True)
    assert nm[0] == n_o and np.array_equal(ms[0], m_o) and n_o > 20


def test_keyframe_matchers_empty_and_degenerate_inputs():
    kf1, kf2, F, epi = make_tri_pair(9, n1=120, n2=150)
    a, b = _to_capi(kf1), _to_capi(kf2)
    fd = kf2.fd
    # KF2 without features / without any FeatureVector entry / F12 == 0 (den == 0 -> CheckDistEpipolarLine false)
    empty = O.KeyFrameData(fd.kps[:0], fd.desc[:0], 640, 480, fd.u_right[:0], SCALE, SIGMA2, kf2.node[:0], kf2.flags[:0])
    nofv = O.KeyFrameData(fd.kps, fd.desc, 640, 480, fd.u_right, SCALE, SIGMA2, kf2.node, np.zeros_like(kf2.flags))
    for other, Fm in [(empty, F), (nofv, F), (kf2, np.zeros(9, np.float32))]:
        nm, ms = capi.search_for_triangulation([(a, _to_capi(other))], [Fm], [epi])
        n_o, m_o = O.search_for_triangulation(kf1, other, Fm, epi)
        assert nm[0] == n_o == 0 and np.array_equal(ms[0], m_o) and (ms[0] == -1).all()
    nm, ms = capi.search_for_triangulation([(_to_capi(empty), b)], [F], [epi])
    assert nm[0] == 0 and len(ms[0]) == 0
    # SearchBySim3 with one empty direction: nothing can agree
    kfa, kfb, q12, d12, i1, q21, d21, i2 = make_sim3_case(31)
    nf, m12 = capi.search_by_sim3(_to_capi(kfa), _to_capi(kfb), q12, d12, i1, q21[:0], d21[:0], i2[:0])
    nf_o, m_o = O.search_by_sim3(kfa, kfb, q12, d12, i1, q21[:0], d21[:0], i2[:0])
    assert nf == nf_o == 0 and np.array_equal(m12, m_o)
    # out-of-range feature index is rejected, not read
    bad = i1.copy(); bad[0] = kfa.n
    with pytest.raises(capi.OrbxError):
        capi.search_by_sim3(_to_capi(kfa), _to_capi(kfb), q12, d12, bad, q21, d21, i2)
    # ComputeDistinctiveDescriptors: no points, and only empty groups
    i0, _ = capi.distinctive_descriptors(np.zeros((0, 32), np.uint8), np.zeros(1, np.int32))
    assert len(i0) == 0
    i3, m3 = capi.distinctive_descriptors(np.zeros((0, 32), np.uint8), np.zeros(4, np.int32))
    assert (i3 == -1).all() and (m3 == -1).all()
