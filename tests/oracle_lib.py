"""ctypes bindings for oracle/liborb_oracle.so (TEST INFRASTRUCTURE ONLY — never imported by the
product package).  Builds the oracle with its Makefile on first use."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
_LIB = None

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28


class Params(C.Structure):
    _fields_ = [("nfeatures", C.c_int32), ("scale_factor", C.c_float), ("nlevels", C.c_int32),
                ("ini_th_fast", C.c_int32), ("min_th_fast", C.c_int32)]


class Frame(C.Structure):
    _fields_ = [("n", C.c_int32), ("keys_un", C.c_void_p), ("descriptors", C.c_void_p),
                ("min_x", C.c_float), ("max_x", C.c_float), ("min_y", C.c_float), ("max_y", C.c_float),
                ("u_right", C.c_void_p), ("scale_factors", C.c_void_p), ("nlevels", C.c_int32)]


class ProjQuery(C.Structure):
    _fields_ = [("u", C.c_float), ("v", C.c_float), ("radius", C.c_float),
                ("min_level", C.c_int32), ("max_level", C.c_int32),
                ("u_right", C.c_float), ("max_right_err", C.c_float), ("angle", C.c_float),
                ("level_lo", C.c_int32), ("level_hi", C.c_int32)]


PROJ_DTYPE = np.dtype([("u", "<f4"), ("v", "<f4"), ("radius", "<f4"), ("min_level", "<i4"),
                       ("max_level", "<i4"), ("u_right", "<f4"), ("max_right_err", "<f4"),
                       ("angle", "<f4"), ("level_lo", "<i4"), ("level_hi", "<i4")])
assert PROJ_DTYPE.itemsize == C.sizeof(ProjQuery)


class StereoInput(C.Structure):
    _fields_ = [("n_left", C.c_int32), ("n_right", C.c_int32),
                ("keys_left", C.c_void_p), ("keys_right", C.c_void_p),
                ("desc_left", C.c_void_p), ("desc_right", C.c_void_p),
                ("nlevels", C.c_int32),
                ("scale_factors", C.c_void_p), ("inv_scale_factors", C.c_void_p),
                ("pyr_left", C.c_void_p), ("pyr_right", C.c_void_p),
                ("level_width", C.c_void_p), ("level_height", C.c_void_p),
                ("pyr_stride_left", C.c_void_p), ("pyr_stride_right", C.c_void_p),
                ("bf", C.c_float), ("fx", C.c_float)]


def build_oracle(force: bool = False) -> str:
    so = os.path.join(ORACLE_DIR, "liborb_oracle.so")
    src = os.path.join(ORACLE_DIR, "orb_oracle.cpp")
    srcs = [src, os.path.join(ORACLE_DIR, "frame_oracle.cpp"), os.path.join(ORACLE_DIR, "bow_oracle.cpp"), os.path.join(ORACLE_DIR, "kfmatch_oracle.cpp")]
    if force or not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(f) for f in srcs):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "liborb_oracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build_oracle())
        L.oracle_fast_atan2.restype = C.c_float
        L.oracle_fast_atan2.argtypes = [C.c_float, C.c_float]
        L.oracle_extractor_create.restype = C.c_void_p
        L.oracle_extractor_create.argtypes = [C.POINTER(Params)]
        L.oracle_extractor_destroy.argtypes = [C.c_void_p]
        L.oracle_extract_many.restype = C.c_int64
        L.oracle_extract_match_many.restype = C.c_int64
        L.oracle_cv_round.argtypes = [C.c_double]
        _LIB = L
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def resize(src: np.ndarray, dw: int, dh: int) -> np.ndarray:
    src = np.ascontiguousarray(src)
    dst = np.empty((dh, dw), np.uint8)
    lib().oracle_resize_linear_u8(_p(src), C.c_int(src.shape[1]), C.c_int(src.shape[0]), C.c_size_t(src.strides[0]),
                                  _p(dst), C.c_int(dw), C.c_int(dh), C.c_size_t(dw))
    return dst


def blur(src: np.ndarray) -> np.ndarray:
    src = np.ascontiguousarray(src)
    dst = np.empty_like(src)
    lib().oracle_gaussian_blur7_u8(_p(src), C.c_int(src.shape[1]), C.c_int(src.shape[0]), C.c_size_t(src.strides[0]),
                                   _p(dst), C.c_size_t(dst.strides[0]))
    return dst


def border101(src: np.ndarray, b: int) -> np.ndarray:
    src = np.ascontiguousarray(src)
    dst = np.empty((src.shape[0] + 2 * b, src.shape[1] + 2 * b), np.uint8)
    lib().oracle_copy_make_border101(_p(src), C.c_int(src.shape[1]), C.c_int(src.shape[0]), C.c_size_t(src.strides[0]),
                                     _p(dst), C.c_size_t(dst.strides[0]), C.c_int(b))
    return dst


def fast(img: np.ndarray, th: int, nms: bool = True) -> np.ndarray:
    """-> float32 [n,3] (x, y, response); img may be a non-contiguous view (row stride honoured)."""
    assert img.strides[1] == 1
    cap = img.shape[0] * img.shape[1]
    out = np.empty((cap, 3), np.float32)
    n = lib().oracle_fast9_16(C.c_void_p(img.ctypes.data), C.c_int(img.shape[1]), C.c_int(img.shape[0]),
                              C.c_size_t(img.strides[0]), C.c_int(th), C.c_int(int(nms)), _p(out), C.c_int(cap))
    return out[:n].copy()


def fast_atan2(y: float, x: float) -> float:
    return float(lib().oracle_fast_atan2(C.c_float(y), C.c_float(x)))


def sincos(x: float, use_libm: bool = False):
    s, c = C.c_float(), C.c_float()
    lib().oracle_sincos(C.c_float(x), C.byref(s), C.byref(c), C.c_int(int(use_libm)))
    return s.value, c.value


class Extractor:
    def __init__(self, nfeatures=1000, scale=1.2, nlevels=8, ini_th=20, min_th=7):
        self.params = Params(nfeatures, scale, nlevels, ini_th, min_th)
        self.h = C.c_void_p(lib().oracle_extractor_create(C.byref(self.params)))
        self.nlevels = nlevels
        self.cap = nfeatures + 64 * nlevels + 64

    def __del__(self):
        if getattr(self, "h", None):
            lib().oracle_extractor_destroy(self.h)
            self.h = None

    def tables(self):
        n = self.nlevels
        sc, isc, s2, is2 = (np.empty(n, np.float32) for _ in range(4))
        fpl = np.empty(n, np.int32)
        umax = np.empty(16, np.int32)
        lib().oracle_extractor_tables(self.h, _p(sc), _p(isc), _p(s2), _p(is2), _p(fpl), _p(umax))
        return dict(scale=sc, inv_scale=isc, sigma2=s2, inv_sigma2=is2, features_per_level=fpl, umax=umax)

    def extract(self, img: np.ndarray, use_libm: bool = False):
        img = np.ascontiguousarray(img)
        kps = np.zeros(self.cap, KP_DTYPE)
        desc = np.zeros((self.cap, 32), np.uint8)
        n = lib().oracle_extract(self.h, _p(img), C.c_int(img.shape[1]), C.c_int(img.shape[0]),
                                 C.c_size_t(img.strides[0]), _p(kps), _p(desc), C.c_int(self.cap),
                                 C.c_int(int(use_libm)))
        assert n >= 0, "oracle capacity"
        return kps[:n].copy(), desc[:n].copy()

    def level(self, level: int, blurred: bool = False) -> np.ndarray:
        w, h = C.c_int(), C.c_int()
        assert lib().oracle_level_size(self.h, C.c_int(level), C.byref(w), C.byref(h)) == 0
        out = np.empty((h.value, w.value), np.uint8)
        assert lib().oracle_get_level(self.h, C.c_int(level), C.c_int(int(blurred)), _p(out), C.c_size_t(w.value)) == 0
        return out

    def candidates(self, level: int) -> np.ndarray:
        n = lib().oracle_get_candidates(self.h, C.c_int(level), None, C.c_int(0))
        out = np.empty((max(n, 1), 3), np.float32)
        lib().oracle_get_candidates(self.h, C.c_int(level), _p(out), C.c_int(n))
        return out[:n]

    def level_keypoints(self, level: int) -> np.ndarray:
        out = np.zeros(self.cap, KP_DTYPE)
        n = lib().oracle_get_level_keypoints(self.h, C.c_int(level), _p(out), C.c_int(self.cap))
        return out[:n].copy()


def distribute_octtree(xyr: np.ndarray, minX, maxX, minY, maxY, N) -> np.ndarray:
    xyr = np.ascontiguousarray(xyr, np.float32)
    n = len(xyr)
    out = np.empty(max(n, 1), np.int32)
    m = lib().oracle_distribute_octtree(_p(xyr), C.c_int(n), C.c_int(minX), C.c_int(maxX), C.c_int(minY),
                                        C.c_int(maxY), C.c_int(N), _p(out), C.c_int(n))
    return out[:m].copy()


def descriptor_distance(a: np.ndarray, b: np.ndarray) -> int:
    a = np.ascontiguousarray(a)
    b = np.ascontiguousarray(b)
    return int(lib().oracle_descriptor_distance(_p(a), _p(b)))


class FrameData:
    """Keeps numpy arrays alive behind an orbm_frame struct."""

    def __init__(self, kps, desc, width, height, u_right=None, scale_factors=None, bounds=None):
        self.kps = np.ascontiguousarray(kps, KP_DTYPE)
        self.desc = np.ascontiguousarray(desc, np.uint8)
        self.u_right = None if u_right is None else np.ascontiguousarray(u_right, np.float32)
        self.sf = None if scale_factors is None else np.ascontiguousarray(scale_factors, np.float32)
        b = bounds if bounds is not None else (0.0, float(width), 0.0, float(height))
        self.bounds = b
        self.c = Frame(len(self.kps), self.kps.ctypes.data, self.desc.ctypes.data, b[0], b[1], b[2], b[3],
                       None if self.u_right is None else self.u_right.ctypes.data,
                       None if self.sf is None else self.sf.ctypes.data,
                       0 if self.sf is None else len(self.sf))


def features_in_area(f: FrameData, x, y, r, min_level=-1, max_level=-1) -> np.ndarray:
    out = np.empty(max(f.c.n, 1), np.int32)
    n = lib().oracle_features_in_area(C.byref(f.c), C.c_float(x), C.c_float(y), C.c_float(r), C.c_int(min_level),
                                      C.c_int(max_level), _p(out), C.c_int(f.c.n))
    return out[:n].copy()


def search_for_initialization(f1: FrameData, f2: FrameData, prev_matched: np.ndarray, window=100, nnratio=0.9,
                              check_ori=True):
    pm = np.ascontiguousarray(prev_matched, np.float32).copy()
    m12 = np.empty(max(f1.c.n, 1), np.int32)
    ncmp = C.c_uint64()
    n = lib().oracle_search_for_initialization(C.byref(f1.c), C.byref(f2.c), _p(pm), _p(m12), C.c_int(window),
                                               C.c_float(nnratio), C.c_int(int(check_ori)), C.byref(ncmp))
    return n, m12[:f1.c.n].copy(), pm, ncmp.value


def search_by_projection(variant, f: FrameData, queries: np.ndarray, qdesc: np.ndarray, blocks, occupied, th_dist,
                         nnratio=0.6, check_ori=True):
    q = np.ascontiguousarray(queries, PROJ_DTYPE)
    qd = np.ascontiguousarray(qdesc, np.uint8)
    nq = len(q)
    occ = np.ascontiguousarray(occupied, np.uint8).copy()
    blk = None if blocks is None else np.ascontiguousarray(blocks, np.uint8)
    assigned = np.empty(max(f.c.n, 1), np.int32)
    bi = np.empty(max(nq, 1), np.int32)
    bd = np.empty(max(nq, 1), np.int32)
    n = lib().oracle_search_by_projection(C.c_int(variant), C.byref(f.c), _p(q), _p(qd), C.c_int(nq), _p(blk),
                                          _p(occ), _p(assigned), _p(bi), _p(bd), C.c_int(th_dist), C.c_float(nnratio),
                                          C.c_int(int(check_ori)))
    return n, assigned[:f.c.n].copy(), bi[:nq].copy(), bd[:nq].copy(), occ


def make_stereo_input(kl, dl, kr, dr, pyr_l, pyr_r, scale, inv_scale, bf, fx):
    """pyr_l / pyr_r: lists of contiguous uint8 level images.  Returns (struct, keepalive)."""
    nl = len(pyr_l)
    kl = np.ascontiguousarray(kl, KP_DTYPE)
    kr = np.ascontiguousarray(kr, KP_DTYPE)
    dl = np.ascontiguousarray(dl, np.uint8)
    dr = np.ascontiguousarray(dr, np.uint8)
    pyr_l = [np.ascontiguousarray(p) for p in pyr_l]
    pyr_r = [np.ascontiguousarray(p) for p in pyr_r]
    pl = (C.c_void_p * nl)(*[p.ctypes.data for p in pyr_l])
    pr = (C.c_void_p * nl)(*[p.ctypes.data for p in pyr_r])
    lw = np.array([p.shape[1] for p in pyr_l], np.int32)
    lh = np.array([p.shape[0] for p in pyr_l], np.int32)
    sl = np.array([p.strides[0] for p in pyr_l], np.int64)
    sr = np.array([p.strides[0] for p in pyr_r], np.int64)
    sc = np.ascontiguousarray(scale, np.float32)
    isc = np.ascontiguousarray(inv_scale, np.float32)
    s = StereoInput(len(kl), len(kr), kl.ctypes.data, kr.ctypes.data, dl.ctypes.data, dr.ctypes.data, nl,
                    sc.ctypes.data, isc.ctypes.data, C.cast(pl, C.c_void_p), C.cast(pr, C.c_void_p),
                    lw.ctypes.data, lh.ctypes.data, sl.ctypes.data, sr.ctypes.data, bf, fx)
    keep = (kl, kr, dl, dr, pyr_l, pyr_r, pl, pr, lw, lh, sl, sr, sc, isc)
    return s, keep


def compute_stereo_matches(kl, dl, kr, dr, pyr_l, pyr_r, scale, inv_scale, bf, fx):
    s, keep = make_stereo_input(kl, dl, kr, dr, pyr_l, pyr_r, scale, inv_scale, bf, fx)
    ur = np.empty(max(len(kl), 1), np.float32)
    dp = np.empty(max(len(kl), 1), np.float32)
    ncmp = C.c_uint64()
    lib().oracle_compute_stereo_matches(C.byref(s), _p(ur), _p(dp), C.byref(ncmp))
    return ur[:len(kl)].copy(), dp[:len(kl)].copy(), ncmp.value


def knn2(queries: np.ndarray, database: np.ndarray, offset=0, nthreads=1):
    q = np.ascontiguousarray(queries, np.uint8)
    d = np.ascontiguousarray(database, np.uint8)
    od = np.empty((len(q), 2), np.uint16)
    oi = np.empty((len(q), 2), np.int32)
    lib().oracle_knn2(_p(q), C.c_int64(len(q)), _p(d), C.c_int64(len(d)), C.c_int32(offset), _p(od), _p(oi),
                      C.c_int(nthreads))
    return od, oi


def extract_many(params: Params, imgs: np.ndarray, nthreads: int) -> int:
    imgs = np.ascontiguousarray(imgs)
    n, h, w = imgs.shape
    return int(lib().oracle_extract_many(C.byref(params), _p(imgs), C.c_size_t(imgs.strides[0]), C.c_int(w),
                                         C.c_int(h), C.c_size_t(imgs.strides[1]), C.c_int(n), C.c_int(nthreads)))


def extract_match_many(params: Params, imgs: np.ndarray, nthreads: int, window=100, nnratio=0.9, check_ori=True):
    """config-2 CPU baseline driver -> (total keypoints, total matches, descriptor compares)."""
    imgs = np.ascontiguousarray(imgs)
    n, h, w = imgs.shape
    tm, nc = C.c_int64(), C.c_uint64()
    tk = lib().oracle_extract_match_many(C.byref(params), _p(imgs), C.c_size_t(imgs.strides[0]), C.c_int(w), C.c_int(h),
                                         C.c_size_t(imgs.strides[1]), C.c_int(n), C.c_int(nthreads), C.c_int(window),
                                         C.c_float(nnratio), C.c_int(int(check_ori)), C.byref(tm), C.byref(nc))
    return int(tk), tm.value, nc.value


# ---- RGB-D frame front (oracle/frame_oracle.cpp) -----------------------------------------------------
def cvt_gray(img: np.ndarray, rgb_order: bool) -> np.ndarray:
    """cv::cvtColor(RGB|BGR|RGBA|BGRA -> GRAY), 8U"""
    img = np.ascontiguousarray(img, np.uint8)
    h, w, ch = img.shape
    out = np.empty((h, w), np.uint8)
    lib().oracle_cvt_gray(_p(img), w, h, C.c_size_t(w * ch), ch, int(rgb_order), _p(out), C.c_size_t(w))
    return out


def depth_to_float(depth: np.ndarray, factor: float) -> np.ndarray:
    depth = np.ascontiguousarray(depth, np.uint16)
    h, w = depth.shape
    out = np.empty((h, w), np.float32)
    lib().oracle_depth_to_float(_p(depth), w, h, C.c_size_t(w), C.c_float(factor), _p(out), C.c_size_t(w))
    return out


def undistort_points(xy: np.ndarray, cam, dist) -> np.ndarray:
    xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    cam = np.asarray(cam, np.float32); dist = np.asarray(dist, np.float32)
    out = np.empty_like(xy)
    lib().oracle_undistort_points(_p(xy), len(xy), _p(cam), _p(dist), _p(out))
    return out


def image_bounds(cols: int, rows: int, cam, dist) -> np.ndarray:
    cam = np.asarray(cam, np.float32); dist = np.asarray(dist, np.float32)
    out = np.empty(4, np.float32)
    lib().oracle_image_bounds(cols, rows, _p(cam), _p(dist), _p(out))
    return out


def stereo_from_rgbd(keys_xy: np.ndarray, keys_un_xy: np.ndarray, depth: np.ndarray, bf: float):
    keys_xy = np.ascontiguousarray(keys_xy, np.float32).reshape(-1, 2)
    keys_un_xy = np.ascontiguousarray(keys_un_xy, np.float32).reshape(-1, 2)
    depth = np.ascontiguousarray(depth, np.float32)
    h, w = depth.shape
    n = len(keys_xy)
    ur = np.empty(n, np.float32); z = np.empty(n, np.float32)
    lib().oracle_stereo_from_rgbd(_p(keys_xy), _p(keys_un_xy), n, _p(depth), w, h, C.c_size_t(w), C.c_float(bf), _p(ur), _p(z))
    return ur, z


# ---- bag-of-words path (oracle/bow_oracle.cpp) ----------------------------------------------------------
def make_vocabulary(k=8, depth=3, seed=0, stop_fraction=0.05):
    """Synthetic DBoW2-shaped vocabulary (the reference's ORBvoc is not shipped): a k-ary tree of the given depth,
    node descriptors = the parent's with ~40 random bits flipped, leaves numbered as words with positive idf-like
    weights (a few 0 = stopped words).  Returned flattened, node 0 = root."""
    rng = np.random.default_rng(seed)
    desc = [np.zeros(32, np.uint8)]
    child_begin, child_count, children, level = [0], [0], [], [0]
    frontier = [0]
    for lv in range(1, depth + 1):
        nxt = []
        for p in frontier:
            child_begin[p] = len(children); child_count[p] = k
            for _ in range(k):
                nid = len(desc)
                bits = np.unpackbits(desc[p] if p else rng.integers(0, 256, 32, dtype=np.uint8))
                flip = rng.choice(256, 40, replace=False)
                bits[flip] ^= 1
                desc.append(np.packbits(bits))
                child_begin.append(0); child_count.append(0); level.append(lv)
                children.append(nid); nxt.append(nid)
        frontier = nxt
    n = len(desc)
    word = np.zeros(n, np.uint32); weight = np.zeros(n, np.float64)
    for w, nid in enumerate(frontier):
        word[nid] = w
        weight[nid] = 0.0 if rng.random() < stop_fraction else float(rng.uniform(0.5, 9.0))
    return dict(nnodes=n, depth=depth, k=k, child_begin=np.array(child_begin, np.int32), child_count=np.array(child_count, np.int32),
                children=np.array(children, np.int32), desc=np.ascontiguousarray(np.stack(desc)), word=word, weight=weight)


def bow_transform(voc, feats: np.ndarray, levelsup: int):
    feats = np.ascontiguousarray(feats, np.uint8)
    n = len(feats)
    wid = np.zeros(n, np.uint32); w = np.zeros(n, np.float64); nid = np.zeros(n, np.uint32)
    lib().oracle_bow_transform(voc["nnodes"], _p(voc["child_begin"]), _p(voc["child_count"]), _p(voc["children"]), _p(voc["desc"]),
                               _p(voc["word"]), _p(voc["weight"]), voc["depth"], _p(feats), n, levelsup, _p(wid), _p(w), _p(nid))
    return wid, w, nid


def search_by_bow(variant, descA, angleA, nodeA, flagsA, descB, angleB, nodeB, flagsB, nnratio=0.7, check_ori=True):
    descA = np.ascontiguousarray(descA, np.uint8); descB = np.ascontiguousarray(descB, np.uint8)
    angleA = np.ascontiguousarray(angleA, np.float32); angleB = np.ascontiguousarray(angleB, np.float32)
    nodeA = np.ascontiguousarray(nodeA, np.uint32); nodeB = np.ascontiguousarray(nodeB, np.uint32)
    flagsA = np.ascontiguousarray(flagsA, np.uint8); flagsB = np.ascontiguousarray(flagsB, np.uint8)
    mA = np.empty(len(descA), np.int32); mB = np.empty(len(descB), np.int32)
    L = lib()
    L.oracle_search_by_bow.restype = C.c_int
    n = L.oracle_search_by_bow(int(variant), len(descA), _p(descA), _p(angleA), _p(nodeA), _p(flagsA), len(descB), _p(descB), _p(angleB),
                               _p(nodeB), _p(flagsB), C.c_float(nnratio), int(check_ori), _p(mA), _p(mB))
    return n, mA, mB


def distinctive_descriptors(desc: np.ndarray, offsets: np.ndarray):
    desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    offsets = np.ascontiguousarray(offsets, np.int32)
    n = len(offsets) - 1
    idx = np.zeros(n, np.int32); med = np.zeros(n, np.int32)
    lib().oracle_distinctive(_p(desc), _p(offsets), n, _p(idx), _p(med))
    return idx, med


class KeyFrameStruct(C.Structure):
    _fields_ = [("frame", Frame), ("level_sigma2", C.c_void_p), ("inv_level_sigma2", C.c_void_p),
                ("node_id", C.c_void_p), ("flags", C.c_void_p)]


class KeyFrameData:
    """numpy arrays behind an orbm_keyframe."""

    def __init__(self, kps, desc, width, height, u_right=None, scale_factors=None, level_sigma2=None, node_id=None,
                 flags=None, bounds=None):
        self.fd = FrameData(kps, desc, width, height, u_right, scale_factors, bounds)
        self.n = self.fd.c.n
        self.s2 = None if level_sigma2 is None else np.ascontiguousarray(level_sigma2, np.float32)
        self.is2 = None if self.s2 is None else (np.float32(1.0) / self.s2).astype(np.float32)
        self.node = None if node_id is None else np.ascontiguousarray(node_id, np.uint32)
        self.flags = None if flags is None else np.ascontiguousarray(flags, np.uint8)
        self.c = KeyFrameStruct(self.fd.c, *[None if a is None else a.ctypes.data for a in (self.s2, self.is2, self.node, self.flags)])


def fuse(variant, kf: KeyFrameData, queries, qdesc):
    q = np.ascontiguousarray(queries, PROJ_DTYPE)
    qd = np.ascontiguousarray(qdesc, np.uint8)
    bi = np.full(max(len(q), 1), -1, np.int32); bd = np.full(max(len(q), 1), 256, np.int32)
    lib().oracle_fuse(C.c_int(variant), C.byref(kf.c), _p(q), _p(qd), C.c_int(len(q)), _p(bi), _p(bd))
    return bi[:len(q)].copy(), bd[:len(q)].copy()


def search_by_sim3(kf1: KeyFrameData, kf2: KeyFrameData, q12, d12, index1, q21, d21, index2):
    q12 = np.ascontiguousarray(q12, PROJ_DTYPE); q21 = np.ascontiguousarray(q21, PROJ_DTYPE)
    d12 = np.ascontiguousarray(d12, np.uint8); d21 = np.ascontiguousarray(d21, np.uint8)
    i1 = np.ascontiguousarray(index1, np.int32); i2 = np.ascontiguousarray(index2, np.int32)
    m12 = np.full(max(kf1.n, 1), -1, np.int32)
    nf = lib().oracle_search_by_sim3(C.byref(kf1.c), C.byref(kf2.c), _p(q12), _p(d12), _p(i1), C.c_int(len(q12)), _p(q21),
                                     _p(d21), _p(i2), C.c_int(len(q21)), _p(m12))
    return nf, m12[:kf1.n].copy()


def search_for_triangulation(kf1: KeyFrameData, kf2: KeyFrameData, f12, epipole, only_stereo=False, check_ori=True):
    F = np.ascontiguousarray(f12, np.float32).reshape(9)
    m = np.full(max(kf1.n, 1), -1, np.int32)
    nm = lib().oracle_search_for_triangulation(C.byref(kf1.c), C.byref(kf2.c), _p(F), C.c_float(epipole[0]), C.c_float(epipole[1]),
                                               C.c_int(int(only_stereo)), C.c_int(int(check_ori)), _p(m))
    return nm, m[:kf1.n].copy()


CLOUD_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("z", "<f4"), ("b", "u1"), ("g", "u1"), ("r", "u1"), ("a", "u1")])


def generate_point_cloud(depth, bgr, fx, fy, cx, cy, pose_inv=None, step=3):
    depth = np.ascontiguousarray(depth, np.float32)
    H, W = depth.shape
    if bgr is not None:
        bgr = np.ascontiguousarray(bgr, np.uint8).reshape(H, W, 3)
    if pose_inv is not None:
        pose_inv = np.ascontiguousarray(pose_inv, np.float64).reshape(16)
    out = np.zeros((-(-W // step)) * (-(-H // step)), CLOUD_DTYPE)
    n = lib().oracle_generate_point_cloud(_p(depth), C.c_size_t(W), _p(bgr), C.c_size_t(W * 3), C.c_int(W), C.c_int(H), C.c_int(step),
                                          C.c_float(fx), C.c_float(fy), C.c_float(cx), C.c_float(cy), _p(pose_inv), _p(out))
    return out[:n].copy()


def write_keyframe_features(kps, desc, mp_index=None):
    kps = np.ascontiguousarray(kps, KP_DTYPE); desc = np.ascontiguousarray(desc, np.uint8)
    n = len(kps)
    out = np.zeros((max(n, 1), 64), np.uint8)
    has = None if mp_index is None else np.ascontiguousarray(np.asarray(mp_index) != np.uint64(0xFFFFFFFFFFFFFFFF), np.uint8)
    mp = None if mp_index is None else np.ascontiguousarray(mp_index, np.uint64)
    lib().oracle_write_keyframe_features.restype = C.c_size_t
    nb = lib().oracle_write_keyframe_features(_p(kps), _p(desc), _p(mp), _p(has), C.c_int(n), _p(out))
    assert nb == n * 64
    return out[:n].copy()


def voxel_grid(points: np.ndarray, leaf: float):
    pts = np.ascontiguousarray(points, CLOUD_DTYPE)
    out = np.zeros(max(len(pts), 1), CLOUD_DTYPE)
    n = lib().oracle_voxel_grid(_p(pts), C.c_int(len(pts)), C.c_float(leaf), _p(out), C.c_int(len(out)))
    assert n >= 0, n
    return out[:n].copy()


def statistical_outlier_removal(points: np.ndarray, mean_k: int = 50, stddev_mul: float = 1.0):
    """-> (kept points, mean distance per input point, [threshold, mean, stddev, valid])"""
    pts = np.ascontiguousarray(points, CLOUD_DTYPE)
    out = np.zeros(max(len(pts), 1), CLOUD_DTYPE)
    md = np.zeros(max(len(pts), 1), np.float32)
    st = np.zeros(4, np.float64)
    f = lib().oracle_statistical_outlier_removal
    f.restype = C.c_longlong
    n = f(_p(pts), C.c_longlong(len(pts)), C.c_int(mean_k), C.c_double(stddev_mul), _p(out), _p(md), _p(st))
    return out[:n].copy(), md[:len(pts)].copy(), st
