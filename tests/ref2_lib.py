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
