"""ctypes bindings for oracle/_ref/libref_matchers.so — the UNMODIFIED reference src/ORBmatcher.cc, Frame.cc,
KeyFrame.cc, MapPoint.cc, Map.cc, KeyFrameDatabase.cc, ORBextractor.cc and the vendored DBoW2 compiled over
oracle/ref_shim2 and driven by oracle/ref_shim2/ref_driver2.cpp (TEST INFRASTRUCTURE ONLY).  The library exists only
where oracle/Makefile could see /root/reference (or where a prebuilt copy travelled with gpurun)."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

import oracle_lib as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF2_SO = os.path.join(ROOT, "oracle", "_ref", "libref_matchers.so")
_LIB = None


def available() -> bool:
    return os.path.exists(REF2_SO)


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(REF2_SO)
    return _LIB


def _p(a):
    if a is None:
        return None
    return C.c_void_p(a.ctypes.data)


def _f32(a):
    return np.ascontiguousarray(a, np.float32)


def _i32(a):
    return np.ascontiguousarray(a, np.int32)


def _u8(a):
    return np.ascontiguousarray(a, np.uint8)


def descriptor_distance(a, b) -> int:
    a, b = _u8(a), _u8(b)
    return int(lib().ref2_descriptor_distance(_p(a), _p(b)))


def features_in_area(f: O.FrameData, x, y, r, min_level=-1, max_level=-1) -> np.ndarray:
    out = np.empty(max(f.c.n, 1), np.int32)
    n = lib().ref2_features_in_area(C.byref(f.c), C.c_float(x), C.c_float(y), C.c_float(r), C.c_int(min_level),
                                    C.c_int(max_level), _p(out), C.c_int(f.c.n))
    return out[:n].copy()


def search_for_initialization(f1: O.FrameData, f2: O.FrameData, prev_matched, window=100, nnratio=0.9, check_ori=True):
    pm = _f32(prev_matched).copy()
    m12 = np.empty(max(f1.c.n, 1), np.int32)
    n = lib().ref2_search_for_initialization(C.byref(f1.c), C.byref(f2.c), _p(pm), _p(m12), C.c_int(window),
                                             C.c_float(nnratio), C.c_int(int(check_ori)))
    return n, m12[:f1.c.n].copy(), pm


def compute_stereo_matches(kl, dl, kr, dr, pyr_l, pyr_r, scale, inv_scale, bf, fx):
    s, keep = O.make_stereo_input(kl, dl, kr, dr, pyr_l, pyr_r, scale, inv_scale, bf, fx)
    ur = np.empty(max(len(kl), 1), np.float32)
    dp = np.empty(max(len(kl), 1), np.float32)
    rc = lib().ref2_compute_stereo_matches(C.byref(s), _p(ur), _p(dp))
    assert rc == 0, "scale tables handed to the reference differ from the ones its own ORBextractor computes"
    return ur[:len(kl)].copy(), dp[:len(kl)].copy()


def proj_localmap(f: O.FrameData, slot_state, proj_x, proj_y, proj_xr, level, view_cos, in_view, qdesc, nobs, th,
                  nnratio, check_ori):
    nq = len(proj_x)
    a = np.empty(max(f.c.n, 1), np.int32)
    args = [_u8(slot_state), _f32(proj_x), _f32(proj_y), _f32(proj_xr), _i32(level), _f32(view_cos), _u8(in_view),
            _u8(qdesc), _i32(nobs)]
    n = lib().ref2_proj_localmap(C.byref(f.c), _p(args[0]), C.c_int(nq), _p(args[1]), _p(args[2]), _p(args[3]),
                                 _p(args[4]), _p(args[5]), _p(args[6]), _p(args[7]), _p(args[8]), C.c_float(th),
                                 C.c_float(nnratio), C.c_int(int(check_ori)), _p(a))
    return n, a[:f.c.n].copy()


def proj_lastframe(cur: O.FrameData, slot_state, bf, u, v, octave, angle, qdesc, nobs, outlier, last_tz, th, mono,
                   nnratio, check_ori):
    nq = len(u)
    a = np.empty(max(cur.c.n, 1), np.int32)
    args = [_u8(slot_state), _f32(u), _f32(v), _i32(octave), _f32(angle), _u8(qdesc), _i32(nobs), _u8(outlier)]
    n = lib().ref2_proj_lastframe(C.byref(cur.c), _p(args[0]), C.c_float(bf), C.c_int(nq), _p(args[1]), _p(args[2]),
                                  _p(args[3]), _p(args[4]), _p(args[5]), _p(args[6]), _p(args[7]), C.c_float(last_tz),
                                  C.c_float(th), C.c_int(int(mono)), C.c_float(nnratio), C.c_int(int(check_ori)), _p(a))
    return n, a[:cur.c.n].copy()


def proj_keyframe(cur: O.FrameData, slot_state, u, v, max_dist, angle, qdesc, already_found, th, orb_dist, nnratio,
                  check_ori):
    nq = len(u)
    a = np.empty(max(cur.c.n, 1), np.int32)
    lv = np.empty(max(nq, 1), np.int32)
    args = [_u8(slot_state), _f32(u), _f32(v), _f32(max_dist), _f32(angle), _u8(qdesc), _u8(already_found)]
    n = lib().ref2_proj_keyframe(C.byref(cur.c), _p(args[0]), C.c_int(nq), _p(args[1]), _p(args[2]), _p(args[3]),
                                 _p(args[4]), _p(args[5]), _p(args[6]), C.c_float(th), C.c_int(orb_dist),
                                 C.c_float(nnratio), C.c_int(int(check_ori)), _p(a), _p(lv))
    return n, a[:cur.c.n].copy(), lv[:nq].copy()


def proj_sim3(kf: O.FrameData, matched_state, u, v, max_dist, qdesc, th):
    nq = len(u)
    a = np.empty(max(kf.c.n, 1), np.int32)
    lv = np.empty(max(nq, 1), np.int32)
    args = [_u8(matched_state), _f32(u), _f32(v), _f32(max_dist), _u8(qdesc)]
    n = lib().ref2_proj_sim3(C.byref(kf.c), _p(args[0]), C.c_int(nq), _p(args[1]), _p(args[2]), _p(args[3]),
                             _p(args[4]), C.c_int(th), _p(a), _p(lv))
    return n, a[:kf.c.n].copy(), lv[:nq].copy()


# ---- SURVEY §8(f) rows ------------------------------------------------------------------------------------------
class RefVocabulary:
    """ORBVocabulary built by the reference's own loadFromTextFile from the flat synthetic tree of oracle_lib.make_vocabulary"""

    def __init__(self, voc, tmp_path="/tmp/orbx_ref2_voc.txt"):
        n = voc["nnodes"]
        parent = np.zeros(n, np.int32)
        for p in range(n):
            b, c = int(voc["child_begin"][p]), int(voc["child_count"][p])
            parent[voc["children"][b:b + c]] = p
        leaf = (voc["child_count"] == 0).astype(np.uint8)
        L = lib()
        L.ref2_voc_create.restype = C.c_void_p
        self.keep = (parent, leaf, _u8(voc["desc"]), np.ascontiguousarray(voc["weight"], np.float64))
        self.h = C.c_void_p(L.ref2_voc_create(C.c_int(voc["k"]), C.c_int(voc["depth"]), C.c_int(n), _p(parent), _p(leaf), _p(self.keep[2]),
                                             _p(self.keep[3]), tmp_path.encode()))
        assert self.h.value, "the reference's loadFromTextFile rejected the vocabulary"

    def __del__(self):
        if getattr(self, "h", None) and self.h.value:
            lib().ref2_voc_destroy(self.h)
            self.h = None

    def transform(self, feats, levelsup):
        feats = _u8(feats)
        n = len(feats)
        w = np.zeros(n, np.uint32); val = np.zeros(n, np.float64); nid = np.zeros(n, np.uint32)
        lib().ref2_voc_transform(self.h, _p(feats), C.c_int(n), C.c_int(levelsup), _p(w), _p(val), _p(nid))
        return w, val, nid

    def transform_all(self, feats, levelsup):
        feats = _u8(feats)
        n = len(feats)
        bw = np.zeros(n, np.uint32); bv = np.zeros(n, np.float64); fn = np.zeros(n, np.uint32); ff = np.zeros(n, np.uint32)
        nfv = C.c_int()
        nb = lib().ref2_voc_transform_all(self.h, _p(feats), C.c_int(n), C.c_int(levelsup), _p(bw), _p(bv), _p(fn), _p(ff), C.byref(nfv))
        return bw[:nb].copy(), bv[:nb].copy(), fn[:nfv.value].copy(), ff[:nfv.value].copy()


def search_by_bow(variant, fa: O.FrameData, node_a, flags_a, fb: O.FrameData, node_b, flags_b, nnratio, check_ori):
    ma = np.empty(max(fa.c.n, 1), np.int32); mb = np.empty(max(fb.c.n, 1), np.int32)
    args = [np.ascontiguousarray(node_a, np.uint32), _u8(flags_a), np.ascontiguousarray(node_b, np.uint32), _u8(flags_b)]
    n = lib().ref2_search_by_bow(C.c_int(variant), C.byref(fa.c), _p(args[0]), _p(args[1]), C.byref(fb.c), _p(args[2]), _p(args[3]),
                                 C.c_float(nnratio), C.c_int(int(check_ori)), _p(ma), _p(mb))
    return n, ma[:fa.c.n].copy(), mb[:fb.c.n].copy()


def search_for_triangulation(f1: O.FrameData, node1, flags1, f2: O.FrameData, node2, flags2, f12, epipole, only_stereo, check_ori):
    m = np.empty(max(f1.c.n, 1), np.int32)
    F = _f32(f12).reshape(9)
    args = [np.ascontiguousarray(node1, np.uint32), _u8(flags1), np.ascontiguousarray(node2, np.uint32), _u8(flags2)]
    n = lib().ref2_search_for_triangulation(C.byref(f1.c), _p(args[0]), _p(args[1]), C.byref(f2.c), _p(args[2]), _p(args[3]), _p(F),
                                            C.c_float(epipole[0]), C.c_float(epipole[1]), C.c_int(int(only_stereo)),
                                            C.c_int(int(check_ori)), _p(m))
    return n, m[:f1.c.n].copy()


def fuse(variant, kf: O.FrameData, bf, u, v, max_dist, qdesc, nobs, th):
    nq = len(u)
    bi = np.empty(max(nq, 1), np.int32); lv = np.empty(max(nq, 1), np.int32)
    args = [_f32(u), _f32(v), _f32(max_dist), _u8(qdesc), _i32(nobs)]
    n = lib().ref2_fuse(C.c_int(variant), C.byref(kf.c), C.c_float(bf), C.c_int(nq), _p(args[0]), _p(args[1]), _p(args[2]), _p(args[3]),
                        _p(args[4]), C.c_float(th), _p(bi), _p(lv))
    return n, bi[:nq].copy(), lv[:nq].copy()


def search_by_sim3(f1: O.FrameData, f2: O.FrameData, has1, u1, v1, maxd1, has2, u2, v2, maxd2, pre12, th):
    m = np.empty(max(f1.c.n, 1), np.int32)
    l1 = np.full(max(f1.c.n, 1), -1, np.int32); l2 = np.full(max(f2.c.n, 1), -1, np.int32)
    args = [_u8(has1), _f32(u1), _f32(v1), _f32(maxd1), _u8(has2), _f32(u2), _f32(v2), _f32(maxd2), _i32(pre12)]
    n = lib().ref2_search_by_sim3(C.byref(f1.c), C.byref(f2.c), *[_p(a) for a in args], C.c_float(th), _p(m), _p(l1), _p(l2))
    return n, m[:f1.c.n].copy(), l1[:f1.c.n].copy(), l2[:f2.c.n].copy()


def distinctive(desc):
    desc = _u8(desc)
    out = np.zeros(32, np.uint8)
    lib().ref2_distinctive(_p(desc), C.c_int(len(desc)), _p(out))
    return out
