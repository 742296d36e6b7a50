"""ctypes view of liborbx.so (the C ABI in include/orbx.h).

This is plumbing for tests and bench.py: the product is the shared library and the C++ wrapper
classes in orb_slam2_map_b200/host/.  There is no CPU fallback — if the library is missing or no
CUDA device is present, calls raise.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG_DIR, "liborbx.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
PROJ_DTYPE = np.dtype([("u", "<f4"), ("v", "<f4"), ("radius", "<f4"), ("min_level", "<i4"),
                       ("max_level", "<i4"), ("u_right", "<f4"), ("max_right_err", "<f4"),
                       ("angle", "<f4"), ("level_lo", "<i4"), ("level_hi", "<i4")])

PROJ_LOCALMAP, PROJ_LASTFRAME, PROJ_KEYFRAME, PROJ_SIM3 = 0, 1, 2, 3


class OrbxError(RuntimeError):
    pass


class Params(C.Structure):
    _fields_ = [("nfeatures", C.c_int32), ("scale_factor", C.c_float), ("nlevels", C.c_int32),
                ("ini_th_fast", C.c_int32), ("min_th_fast", C.c_int32)]


class FrameStruct(C.Structure):
    _fields_ = [("n", C.c_int32), ("keys_un", C.c_void_p), ("descriptors", C.c_void_p),
                ("min_x", C.c_float), ("max_x", C.c_float), ("min_y", C.c_float), ("max_y", C.c_float),
                ("u_right", C.c_void_p), ("scale_factors", C.c_void_p), ("nlevels", C.c_int32)]


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


STAGES = ["resize", "blur", "fast", "octree", "describe", "grid", "search_init", "stereo"]

EXPORTS = [
    "orbx_set_profiling", "orbx_get_profile",
    "orbx_last_error", "orbx_device_count", "orbx_create", "orbx_destroy", "orbx_sync", "orbx_get_tables",
    "orbx_keypoint_capacity", "orbx_reserve_shape", "orbx_extract", "orbx_extract_batch", "orbx_extract_batch_to_host", "orbx_download", "orbx_get_level",
    "orbx_get_candidates", "orbx_search_init_pairs", "orbx_extract_stereo_batch",
    "orbm_descriptor_distance", "orbm_search_for_initialization", "orbm_search_by_projection",
    "orbm_compute_stereo_matches", "orbm_knn2_device", "orbm_knn2_device_noalloc", "orbm_statistical_outlier_removal", "orbm_knn2_merge_device", "orbm_knn2_set_variant",
    "orbm_measure_int_rates", "orbm_nccl_unique_id", "orbm_nccl_comm_create", "orbm_nccl_comm_destroy", "orbm_knn2_sharded",
    "orbx_cvt_gray", "orbx_depth_to_float", "orbx_undistort_points", "orbx_image_bounds", "orbx_frames_rgbd", "orbx_frame_rgbd",
    "orbx_download_frame_front",
    "orbm_vocabulary_upload", "orbm_vocabulary_free", "orbm_bow_transform", "orbm_search_by_bow", "orbm_distinctive_descriptors",
    "orbm_search_for_triangulation", "orbm_fuse", "orbm_search_by_sim3", "orbm_generate_point_clouds", "orbm_voxel_grid", "orbx_pack_map_records", "orbx_extract_match_to_host",
]


class Camera(C.Structure):
    """orbx_camera: mK, mDistCoef (CV_32F) and mbf"""
    _fields_ = [("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float),
                ("k1", C.c_float), ("k2", C.c_float), ("p1", C.c_float), ("p2", C.c_float), ("k3", C.c_float),
                ("bf", C.c_float)]

_LIB = None


def build(verbose: bool = False) -> str:
    """Compile liborbx.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    out = None if verbose else subprocess.DEVNULL
    subprocess.check_call(["make", "-C", os.path.join(PKG_DIR, "csrc"), "-j4"], stdout=out)
    return LIB_PATH


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise OrbxError(f"{LIB_PATH} is missing: build it with `make -C orb_slam2_map_b200/csrc` "
                            "(there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        L.orbx_last_error.restype = C.c_char_p
        L.orbx_create.argtypes = [C.POINTER(Params), C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                  C.POINTER(C.c_void_p)]
        L.orbx_destroy.argtypes = [C.c_void_p]
        L.orbx_sync.argtypes = [C.c_void_p]
        L.orbx_set_profiling.argtypes = [C.c_void_p, C.c_int]
        L.orbx_get_profile.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.orbx_get_tables.argtypes = [C.c_void_p] + [C.c_void_p] * 5
        L.orbx_keypoint_capacity.argtypes = [C.c_void_p]
        L.orbx_reserve_shape.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.orbx_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_void_p, C.c_void_p,
                                   C.c_int, C.POINTER(C.c_int)]
        L.orbx_extract_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_int, C.c_int, C.c_int,
                                         C.c_int, C.c_int]
        L.orbx_extract_batch_to_host.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_int, C.c_int, C.c_int,
                                                 C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.orbx_download.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.orbx_get_level.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t,
                                     C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.orbx_get_candidates.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_int)]
        L.orbx_search_init_pairs.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, C.c_void_p,
                                             C.c_void_p, C.c_int, C.POINTER(C.c_uint64)]
        L.orbx_extract_stereo_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_int,
                                                C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_void_p,
                                                C.c_void_p, C.c_int, C.POINTER(C.c_uint64)]
        L.orbm_descriptor_distance.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.orbm_search_for_initialization.argtypes = [C.POINTER(FrameStruct), C.POINTER(FrameStruct), C.c_void_p,
                                                     C.c_void_p, C.c_int, C.c_float, C.c_int, C.POINTER(C.c_int32)]
        L.orbm_search_by_projection.argtypes = [C.c_int, C.POINTER(FrameStruct), C.c_void_p, C.c_void_p, C.c_int,
                                                C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                                C.c_float, C.c_int, C.POINTER(C.c_int32)]
        L.orbm_compute_stereo_matches.argtypes = [C.POINTER(StereoInput), C.c_void_p, C.c_void_p,
                                                  C.POINTER(C.c_uint64)]
        L.orbm_knn2_device.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p,
                                       C.c_void_p, C.c_void_p]
        L.orbm_knn2_device_noalloc.argtypes = L.orbm_knn2_device.argtypes
        L.orbm_knn2_merge_device.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int64, C.c_void_p, C.c_void_p,
                                             C.c_void_p]
        L.orbm_nccl_comm_create.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
        L.orbm_nccl_comm_destroy.argtypes = [C.c_void_p]
        L.orbm_knn2_sharded.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int32, C.c_int64,
                                        C.c_void_p, C.c_void_p, C.c_void_p]
        _LIB = L
    return _LIB


def check(rc: int):
    if rc != 0:
        raise OrbxError(f"liborbx error {rc}: {lib().orbx_last_error().decode()}")


def _p(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data_as(C.c_void_p)
    return C.c_void_p(int(a))


class Extractor:
    """Owns one orbx_handle.  Mirrors ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)
    (reference include/ORBextractor.h:51) plus the batch / device-resident entry points."""

    def __init__(self, nfeatures=1000, scale=1.2, nlevels=8, ini_th=20, min_th=7, max_width=640, max_height=480,
                 chunk_frames=1, result_slots=2, device=0, stream=None):
        self.params = Params(nfeatures, scale, nlevels, ini_th, min_th)
        self.h = C.c_void_p()
        self.nlevels = nlevels
        check(lib().orbx_create(C.byref(self.params), max_width, max_height, chunk_frames, result_slots, device,
                                C.c_void_p(stream) if stream else None, C.byref(self.h)))
        self.result_slots = result_slots

    @property
    def cap(self):
        """per-frame keypoint capacity (grows when a shape with more quadtree roots is reserved / seen)"""
        return lib().orbx_keypoint_capacity(self.h)

    def reserve_shape(self, width, height):
        """orbx_reserve_shape: lay out the workspace for this shape, -> keypoint capacity output buffers need."""
        rc = lib().orbx_reserve_shape(self.h, width, height)
        check(min(rc, 0))
        return rc

    def close(self):
        if getattr(self, "h", None) and self.h.value:
            lib().orbx_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def sync(self):
        check(lib().orbx_sync(self.h))

    def set_profiling(self, on: bool):
        check(lib().orbx_set_profiling(self.h, int(on)))

    def profile(self):
        """-> ({stage: ms}, {stage: launches}) accumulated since the previous call."""
        ms = np.zeros(len(STAGES), np.float64)
        ln = np.zeros(len(STAGES), np.int64)
        check(lib().orbx_get_profile(self.h, _p(ms), _p(ln)))
        return dict(zip(STAGES, ms.tolist())), dict(zip(STAGES, ln.tolist()))

    def tables(self):
        n = self.nlevels
        sc, isc, s2, is2 = (np.empty(n, np.float32) for _ in range(4))
        fpl = np.empty(n, np.int32)
        check(lib().orbx_get_tables(self.h, _p(sc), _p(isc), _p(s2), _p(is2), _p(fpl)))
        return dict(scale=sc, inv_scale=isc, sigma2=s2, inv_sigma2=is2, features_per_level=fpl)

    def extract(self, img: np.ndarray):
        """ORBextractor::operator() on one host image -> (keypoints[KP_DTYPE], descriptors[n,32])."""
        img = np.ascontiguousarray(img, np.uint8)
        cap = self.reserve_shape(img.shape[1], img.shape[0]) if img.size else self.cap
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = C.c_int(0)
        check(lib().orbx_extract(self.h, _p(img), img.shape[1], img.shape[0], img.strides[0], _p(kps), _p(desc),
                                 cap, C.byref(n)))
        return kps[:n.value].copy(), desc[:n.value].copy()

    def extract_batch(self, images, nframes=None, width=None, height=None, frame_stride=None, row_stride=None,
                      on_device=False, first_slot=0):
        """images: numpy [n,h,w] (host) or a raw device pointer (int) with explicit geometry."""
        if isinstance(images, np.ndarray):
            assert images.ndim == 3 and images.dtype == np.uint8 and images.strides[2] == 1
            nframes, height, width = images.shape
            frame_stride, row_stride = images.strides[0], images.strides[1]
            ptr = images.ctypes.data
        else:
            ptr = int(images)
        check(lib().orbx_extract_batch(self.h, C.c_void_p(ptr), frame_stride, row_stride, width, height, nframes,
                                       int(on_device), first_slot))

    def extract_batch_to_host(self, images: np.ndarray, kps: np.ndarray, desc: np.ndarray, counts: np.ndarray,
                              first_slot=0):
        """orbx_extract_batch_to_host on host images [n,h,w] with caller-owned (pinned) outputs."""
        assert images.ndim == 3 and images.dtype == np.uint8 and images.strides[2] == 1
        n, h, w = images.shape
        check(lib().orbx_extract_batch_to_host(self.h, C.c_void_p(images.ctypes.data), images.strides[0],
                                               images.strides[1], w, h, n, 0, first_slot, _p(kps), _p(desc),
                                               _p(counts), self.cap))

    def extract_match_to_host(self, images: np.ndarray, kps, desc, counts, matches12, nmatches, window=100, nnratio=0.9,
                              check_ori=True, first_slot=0):
        """orbx_extract_match_to_host: extraction + consecutive-frame SearchForInitialization in one chunk pipeline."""
        assert images.ndim == 3 and images.dtype == np.uint8 and images.strides[2] == 1
        n, h, w = images.shape
        check(lib().orbx_extract_match_to_host(self.h, C.c_void_p(images.ctypes.data), C.c_size_t(images.strides[0]),
                                               C.c_size_t(images.strides[1]), w, h, n, 0, first_slot, _p(kps), _p(desc), _p(counts),
                                               self.cap, window, C.c_float(nnratio), int(check_ori), _p(matches12), _p(nmatches)))

    def download(self, first_slot, nframes, want_kps=True, want_desc=True, out=None):
        counts = np.zeros(nframes, np.int32)
        if out is not None:
            kps, desc = out
        else:
            kps = np.zeros((nframes, self.cap), KP_DTYPE) if want_kps else None
            desc = np.zeros((nframes, self.cap, 32), np.uint8) if want_desc else None
        check(lib().orbx_download(self.h, first_slot, nframes, _p(kps), _p(desc), _p(counts), self.cap))
        return kps, desc, counts

    def level(self, frame, level, blurred=False, border=0):
        w, h = C.c_int(), C.c_int()
        check(lib().orbx_get_level(self.h, frame, level, int(blurred), border, None, 0, C.byref(w), C.byref(h)))
        out = np.empty((h.value + 2 * border, w.value + 2 * border), np.uint8)
        check(lib().orbx_get_level(self.h, frame, level, int(blurred), border, _p(out), out.strides[0], C.byref(w),
                                   C.byref(h)))
        return out

    def candidates(self, frame, level, cap=1 << 18):
        out = np.empty((cap, 3), np.float32)
        n = C.c_int()
        check(lib().orbx_get_candidates(self.h, frame, level, _p(out), cap, C.byref(n)))
        assert n.value <= cap
        return out[:n.value].copy()

    # ---- RGB-D frame front (orbx_frame.cu) ------------------------------------------------------
    def cvt_gray(self, color: np.ndarray, rgb_order: bool) -> np.ndarray:
        color = np.ascontiguousarray(color, np.uint8)
        if color.ndim == 3:
            color = color[None]
        n, h, w, ch = color.shape
        gray = np.empty((n, h, w), np.uint8)
        check(lib().orbx_cvt_gray(self.h, color.ctypes.data_as(C.c_void_p), ch, int(rgb_order), C.c_size_t(h * w * ch),
                                  C.c_size_t(w * ch), w, h, n, 0, gray.ctypes.data_as(C.c_void_p), C.c_size_t(h * w), C.c_size_t(w)))
        return gray

    def depth_to_float(self, depth: np.ndarray, factor: float) -> np.ndarray:
        depth = np.ascontiguousarray(depth, np.uint16)
        h, w = depth.shape
        out = np.empty((h, w), np.float32)
        check(lib().orbx_depth_to_float(self.h, depth.ctypes.data_as(C.c_void_p), C.c_size_t(w), w, h, C.c_float(factor), 0,
                                        out.ctypes.data_as(C.c_void_p), C.c_size_t(w)))
        return out

    def undistort_points(self, xy: np.ndarray, cam: "Camera") -> np.ndarray:
        xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
        out = np.empty_like(xy)
        check(lib().orbx_undistort_points(self.h, xy.ctypes.data_as(C.c_void_p), len(xy), C.byref(cam), out.ctypes.data_as(C.c_void_p)))
        return out

    def image_bounds(self, cols: int, rows: int, cam: "Camera") -> np.ndarray:
        out = np.empty(4, np.float32)
        check(lib().orbx_image_bounds(self.h, cols, rows, C.byref(cam), out.ctypes.data_as(C.c_void_p)))
        return out

    def frames_rgbd(self, color: np.ndarray, depth: np.ndarray, cam: "Camera", rgb_order=True, depth_factor=1.0, first_slot=0):
        """color [n,h,w,c] or [n,h,w] uint8, depth [n,h,w] uint16 or float32 (host arrays)"""
        color = np.ascontiguousarray(color, np.uint8)
        ch = 1 if color.ndim == 3 else color.shape[3]
        n, h, w = color.shape[:3]
        is16 = depth.dtype == np.uint16
        depth = np.ascontiguousarray(depth, np.uint16 if is16 else np.float32)
        es = 2 if is16 else 4
        check(lib().orbx_frames_rgbd(self.h, color.ctypes.data_as(C.c_void_p), ch, int(rgb_order), C.c_size_t(h * w * ch), C.c_size_t(w * ch),
                                     depth.ctypes.data_as(C.c_void_p), int(is16), C.c_float(depth_factor), C.c_size_t(h * w * es),
                                     C.c_size_t(w * es), w, h, n, 0, C.byref(cam), first_slot))

    def frame_rgbd(self, color: np.ndarray, depth: np.ndarray, cam: "Camera", rgb_order=True, depth_factor=1.0):
        """orbx_frame_rgbd: ONE RGB-D frame (host arrays, any row stride) -> dict(kps, keys_un, desc, u_right, depth, bounds)"""
        assert color.dtype == np.uint8 and color.strides[-1] == 1 and (color.ndim == 2 or color.strides[-2] == color.shape[-1])
        ch = 1 if color.ndim == 2 else color.shape[2]
        h, w = color.shape[:2]
        is16 = depth.dtype == np.uint16
        assert depth.dtype in (np.uint16, np.float32) and depth.strides[1] == depth.itemsize
        cap = self.reserve_shape(w, h)
        kps = np.zeros(cap, KP_DTYPE); kun = np.zeros(cap, KP_DTYPE); desc = np.zeros((cap, 32), np.uint8)
        ur = np.zeros(cap, np.float32); z = np.zeros(cap, np.float32); b = np.zeros(4, np.float32)
        n = C.c_int(0)
        check(lib().orbx_frame_rgbd(self.h, C.c_void_p(color.ctypes.data), ch, int(rgb_order), C.c_size_t(color.strides[0]),
                                    C.c_void_p(depth.ctypes.data), int(is16), C.c_float(depth_factor), C.c_size_t(depth.strides[0]), w, h,
                                    C.byref(cam), _p(kps), _p(kun), _p(desc), _p(ur), _p(z), cap, C.byref(n), _p(b)))
        c = n.value
        return dict(kps=kps[:c].copy(), keys_un=kun[:c].copy(), desc=desc[:c].copy(), u_right=ur[:c].copy(), depth=z[:c].copy(), bounds=b)

    def frames_rgbd_ptr(self, color_ptr, channels, depth_ptr, depth_is_u16, nframes, width, height, cam, rgb_order=True,
                        depth_factor=1.0, on_device=True, first_slot=0):
        """raw-pointer form (device or pinned host memory), dense layouts"""
        es = 2 if depth_is_u16 else 4
        check(lib().orbx_frames_rgbd(self.h, C.c_void_p(color_ptr), channels, int(rgb_order), C.c_size_t(height * width * channels),
                                     C.c_size_t(width * channels), C.c_void_p(depth_ptr), int(depth_is_u16), C.c_float(depth_factor),
                                     C.c_size_t(height * width * es), C.c_size_t(width * es), width, height, nframes, int(on_device),
                                     C.byref(cam), first_slot))

    def download_frame_front(self, first_slot, nframes, out=None):
        """out: (mvKeysUn [nframes, cap] KP_DTYPE, mvuRight [nframes, cap] f32, mvDepth [nframes, cap] f32) to fill (e.g. page-locked)"""
        if out is not None:
            kun, ur, z = out
        else:
            kun = np.zeros((nframes, self.cap), KP_DTYPE)
            ur = np.zeros((nframes, self.cap), np.float32)
            z = np.zeros((nframes, self.cap), np.float32)
        b = np.zeros(4, np.float32)
        check(lib().orbx_download_frame_front(self.h, first_slot, nframes, kun.ctypes.data_as(C.c_void_p), ur.ctypes.data_as(C.c_void_p),
                                              z.ctypes.data_as(C.c_void_p), self.cap, b.ctypes.data_as(C.c_void_p)))
        return kun, ur, z, b

    def pack_map_records(self, slot, mp_index=None):
        """Map::Save per-feature records (64 bytes each) of result slot `slot`"""
        out = np.zeros((self.cap, 64), np.uint8)
        n = C.c_int32()
        mp = None if mp_index is None else np.ascontiguousarray(mp_index, np.uint64)
        check(lib().orbx_pack_map_records(self.h, slot, _p(mp), _p(out), self.cap, C.byref(n)))
        return out[:n.value].copy()

    def search_init_pairs(self, first_slot, npairs, window=100, nnratio=0.9, check_ori=True, fetch=True):
        ncmp = C.c_uint64()
        if fetch:
            m12 = np.full((npairs, self.cap), -1, np.int32)
            nm = np.zeros(npairs, np.int32)
            check(lib().orbx_search_init_pairs(self.h, first_slot, npairs, window, nnratio, int(check_ori), _p(m12),
                                               _p(nm), self.cap, C.byref(ncmp)))
            return m12, nm, ncmp.value
        check(lib().orbx_search_init_pairs(self.h, first_slot, npairs, window, nnratio, int(check_ori), None, None,
                                           self.cap, None))
        return None

    def extract_stereo_batch(self, left, right, bf, fx, nframes=None, width=None, height=None, frame_stride=None,
                             row_stride=None, on_device=False, first_slot=0, fetch=True, out=None):
        """out: optional caller-owned (pinned) float32 arrays (u_right, depth) of shape [npairs, cap]"""
        if isinstance(left, np.ndarray):
            assert left.shape == right.shape and left.strides == right.strides
            nframes, height, width = left.shape
            frame_stride, row_stride = left.strides[0], left.strides[1]
            lp, rp = left.ctypes.data, right.ctypes.data
        else:
            lp, rp = int(left), int(right)
        ncmp = C.c_uint64()
        if fetch:
            if out is not None:
                ur, dp = out
            else:
                ur = np.zeros((nframes, self.cap), np.float32)
                dp = np.zeros((nframes, self.cap), np.float32)
            check(lib().orbx_extract_stereo_batch(self.h, C.c_void_p(lp), C.c_void_p(rp), frame_stride, row_stride,
                                                  width, height, nframes, int(on_device), first_slot, bf, fx, _p(ur),
                                                  _p(dp), self.cap, C.byref(ncmp)))
            return ur, dp, ncmp.value
        check(lib().orbx_extract_stereo_batch(self.h, C.c_void_p(lp), C.c_void_p(rp), frame_stride, row_stride, width,
                                              height, nframes, int(on_device), first_slot, bf, fx, None, None,
                                              self.cap, None))
        return None


class FrameData:
    """numpy arrays behind an orbm_frame."""

    def __init__(self, kps, desc, width, height, u_right=None, scale_factors=None, bounds=None):
        self.kps = np.ascontiguousarray(kps, KP_DTYPE)
        self.desc = np.ascontiguousarray(desc, np.uint8)
        self.u_right = None if u_right is None else np.ascontiguousarray(u_right, np.float32)
        self.sf = None if scale_factors is None else np.ascontiguousarray(scale_factors, np.float32)
        b = bounds if bounds is not None else (0.0, float(width), 0.0, float(height))
        self.c = FrameStruct(len(self.kps), self.kps.ctypes.data, self.desc.ctypes.data, b[0], b[1], b[2], b[3],
                             None if self.u_right is None else self.u_right.ctypes.data,
                             None if self.sf is None else self.sf.ctypes.data,
                             0 if self.sf is None else len(self.sf))


def descriptor_distance(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    a = np.ascontiguousarray(a, np.uint8).reshape(-1, 32)
    b = np.ascontiguousarray(b, np.uint8).reshape(-1, 32)
    out = np.empty(len(a), np.int32)
    check(lib().orbm_descriptor_distance(_p(a), _p(b), len(a), _p(out)))
    return out


def search_for_initialization(f1: FrameData, f2: FrameData, prev_matched, window=100, nnratio=0.9, check_ori=True):
    pm = np.ascontiguousarray(prev_matched, np.float32).copy()
    m12 = np.empty(max(f1.c.n, 1), np.int32)
    nm = C.c_int32()
    check(lib().orbm_search_for_initialization(C.byref(f1.c), C.byref(f2.c), _p(pm), _p(m12), window, nnratio,
                                               int(check_ori), C.byref(nm)))
    return nm.value, m12[:f1.c.n].copy(), pm


def search_by_projection(variant, f: FrameData, queries, qdesc, blocks, occupied, th_dist, nnratio=0.6,
                         check_ori=True):
    q = np.ascontiguousarray(queries, PROJ_DTYPE)
    qd = np.ascontiguousarray(qdesc, np.uint8)
    nq = len(q)
    occ = np.ascontiguousarray(occupied, np.uint8).copy()
    blk = None if blocks is None else np.ascontiguousarray(blocks, np.uint8)
    assigned = np.empty(max(f.c.n, 1), np.int32)
    bi = np.empty(max(nq, 1), np.int32)
    bd = np.empty(max(nq, 1), np.int32)
    nm = C.c_int32()
    check(lib().orbm_search_by_projection(variant, C.byref(f.c), _p(q), _p(qd), nq, _p(blk), _p(occ), _p(assigned),
                                          _p(bi), _p(bd), th_dist, nnratio, int(check_ori), C.byref(nm)))
    return nm.value, assigned[:f.c.n].copy(), bi[:nq].copy(), bd[:nq].copy(), occ


def compute_stereo_matches(kl, dl, kr, dr, pyr_l, pyr_r, scale, inv_scale, bf, fx):
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
    ur = np.empty(max(len(kl), 1), np.float32)
    dp = np.empty(max(len(kl), 1), np.float32)
    ncmp = C.c_uint64()
    check(lib().orbm_compute_stereo_matches(C.byref(s), _p(ur), _p(dp), C.byref(ncmp)))
    return ur[:len(kl)].copy(), dp[:len(kl)].copy(), ncmp.value


def knn2_device(q_ptr, nq, db_ptr, ndb, offset, out_dist_ptr, out_idx_ptr, stream=None):
    check(lib().orbm_knn2_device(C.c_void_p(q_ptr), nq, C.c_void_p(db_ptr), ndb, offset, C.c_void_p(out_dist_ptr),
                                 C.c_void_p(out_idx_ptr), C.c_void_p(stream) if stream else None))


def knn2_device_noalloc(q_ptr, nq, db_ptr, ndb, offset, out_dist_ptr, out_idx_ptr, stream=None):
    """orbm_knn2_device_noalloc: the same search with its scratch in a built-in device array (no cudaMalloc in the call)"""
    check(lib().orbm_knn2_device_noalloc(C.c_void_p(q_ptr), nq, C.c_void_p(db_ptr), ndb, offset, C.c_void_p(out_dist_ptr),
                                         C.c_void_p(out_idx_ptr), C.c_void_p(stream) if stream else None))


def knn2_merge_device(pd_ptr, pi_ptr, nparts, nq, od_ptr, oi_ptr, stream=None):
    check(lib().orbm_knn2_merge_device(C.c_void_p(pd_ptr), C.c_void_p(pi_ptr), nparts, nq, C.c_void_p(od_ptr),
                                       C.c_void_p(oi_ptr), C.c_void_p(stream) if stream else None))


def measure_int_rates():
    """orbm_measure_int_rates -> dict(popc, lop3, add: results/s; knn_mix: register-only compares/s)"""
    r = (C.c_double * 4)()
    check(lib().orbm_measure_int_rates(r))
    return dict(popc=r[0], lop3=r[1], add=r[2], knn_mix=r[3])


def nccl_unique_id() -> bytes:
    """orbm_nccl_unique_id: rank 0 makes it, the other ranks receive its 128 bytes"""
    buf = (C.c_uint8 * 128)()
    check(lib().orbm_nccl_unique_id(buf))
    return bytes(buf)


def nccl_comm_create(uid: bytes, rank: int, nranks: int) -> int:
    """orbm_nccl_comm_create on the current CUDA device -> ncclComm_t as an integer handle"""
    buf = (C.c_uint8 * 128).from_buffer_copy(uid)
    comm = C.c_void_p()
    check(lib().orbm_nccl_comm_create(buf, rank, nranks, C.byref(comm)))
    return comm.value


def nccl_comm_destroy(comm: int):
    check(lib().orbm_nccl_comm_destroy(C.c_void_p(comm)))


def knn2_sharded(comm, q_ptr, nq, shard_ptr, ndb_local, offset, out_dist_ptr, out_idx_ptr, stream=None, query_chunk=0):
    """orbm_knn2_sharded: local scan + ncclAllGather + merge, chunked over the queries; comm None = single GPU"""
    check(lib().orbm_knn2_sharded(C.c_void_p(comm) if comm else None, C.c_void_p(q_ptr), nq, C.c_void_p(shard_ptr), ndb_local,
                                  offset, query_chunk, C.c_void_p(out_dist_ptr), C.c_void_p(out_idx_ptr),
                                  C.c_void_p(stream) if stream else None))


# ---- bag-of-words path (orbx_bow.cu) -------------------------------------------------------------------
class VocabularyStruct(C.Structure):
    _fields_ = [("nnodes", C.c_int), ("depth", C.c_int), ("child_begin", C.c_void_p), ("child_count", C.c_void_p),
                ("children", C.c_void_p), ("descriptors", C.c_void_p), ("word_id", C.c_void_p), ("weight", C.c_void_p)]


class BowFrameStruct(C.Structure):
    _fields_ = [("n", C.c_int), ("descriptors", C.c_void_p), ("angles", C.c_void_p), ("node_id", C.c_void_p), ("flags", C.c_void_p)]


class Vocabulary:
    """Device copy of a flattened DBoW2 vocabulary tree (dict of numpy arrays: child_begin, child_count, children,
    desc, word, weight + nnodes, depth)."""

    def __init__(self, voc: dict, device=0):
        self._keep = {k: np.ascontiguousarray(voc[k], dt) for k, dt in (("child_begin", np.int32), ("child_count", np.int32),
                      ("children", np.int32), ("desc", np.uint8), ("word", np.uint32), ("weight", np.float64))}
        k = self._keep
        st = VocabularyStruct(int(voc["nnodes"]), int(voc["depth"]), k["child_begin"].ctypes.data, k["child_count"].ctypes.data,
                              k["children"].ctypes.data, k["desc"].ctypes.data, k["word"].ctypes.data, k["weight"].ctypes.data)
        self.h = C.c_void_p()
        check(lib().orbm_vocabulary_upload(C.byref(st), device, C.byref(self.h)))

    def transform(self, feats: np.ndarray, levelsup: int):
        feats = np.ascontiguousarray(feats, np.uint8)
        n = len(feats)
        wid = np.zeros(n, np.uint32); w = np.zeros(n, np.float64); nid = np.zeros(n, np.uint32)
        check(lib().orbm_bow_transform(self.h, feats.ctypes.data_as(C.c_void_p), n, levelsup, wid.ctypes.data_as(C.c_void_p),
                                       w.ctypes.data_as(C.c_void_p), nid.ctypes.data_as(C.c_void_p)))
        return wid, w, nid

    def close(self):
        if self.h:
            lib().orbm_vocabulary_free(self.h)
            self.h = None


def search_by_bow(variant, pairs, nnratio=0.7, check_ori=True, device=0):
    """pairs: list of ((descA, angleA, nodeA, flagsA), (descB, angleB, nodeB, flagsB)); returns per pair (n, matchA, matchB)"""
    keep = []
    def frame(t):
        d, a, nd, fl = (np.ascontiguousarray(t[0], np.uint8), np.ascontiguousarray(t[1], np.float32),
                        np.ascontiguousarray(t[2], np.uint32), np.ascontiguousarray(t[3], np.uint8))
        keep.append((d, a, nd, fl))
        return BowFrameStruct(len(d), d.ctypes.data, a.ctypes.data, nd.ctypes.data, fl.ctypes.data)
    A = (BowFrameStruct * len(pairs))(*[frame(p[0]) for p in pairs])
    B = (BowFrameStruct * len(pairs))(*[frame(p[1]) for p in pairs])
    totA = sum(len(p[0][0]) for p in pairs); totB = sum(len(p[1][0]) for p in pairs)
    mA = np.zeros(max(totA, 1), np.int32); mB = np.zeros(max(totB, 1), np.int32); nm = np.zeros(len(pairs), np.int32)
    check(lib().orbm_search_by_bow(int(variant), A, B, len(pairs), C.c_float(nnratio), int(check_ori), device,
                                   mA.ctypes.data_as(C.c_void_p), mB.ctypes.data_as(C.c_void_p), nm.ctypes.data_as(C.c_void_p)))
    out, oa, ob = [], 0, 0
    for i, p in enumerate(pairs):
        na, nb = len(p[0][0]), len(p[1][0])
        out.append((int(nm[i]), mA[oa:oa + na].copy(), mB[ob:ob + nb].copy()))
        oa += na; ob += nb
    return out


def distinctive_descriptors(desc: np.ndarray, offsets: np.ndarray, device=0):
    """MapPoint::ComputeDistinctiveDescriptors for len(offsets)-1 groups of descriptors"""
    desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    offsets = np.ascontiguousarray(offsets, np.int32)
    n = len(offsets) - 1
    idx = np.zeros(n, np.int32); med = np.zeros(n, np.int32)
    check(lib().orbm_distinctive_descriptors(desc.ctypes.data_as(C.c_void_p), offsets.ctypes.data_as(C.c_void_p), n, device,
                                             idx.ctypes.data_as(C.c_void_p), med.ctypes.data_as(C.c_void_p)))
    return idx, med


class KeyFrameStruct(C.Structure):
    _fields_ = [("frame", FrameStruct), ("level_sigma2", C.c_void_p), ("inv_level_sigma2", C.c_void_p),
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


def fuse(variant, kf: KeyFrameData, queries, qdesc, device=0):
    """per-MapPoint search of ORBmatcher::Fuse (variant 0 / 1) or of one SearchBySim3 direction (2)"""
    q = np.ascontiguousarray(queries, PROJ_DTYPE)
    qd = np.ascontiguousarray(qdesc, np.uint8)
    bi = np.full(max(len(q), 1), -1, np.int32); bd = np.full(max(len(q), 1), 256, np.int32)
    check(lib().orbm_fuse(variant, C.byref(kf.c), _p(q), _p(qd), len(q), device, _p(bi), _p(bd)))
    return bi[:len(q)].copy(), bd[:len(q)].copy()


def search_by_sim3(kf1: KeyFrameData, kf2: KeyFrameData, q12, d12, index1, q21, d21, index2, device=0):
    q12 = np.ascontiguousarray(q12, PROJ_DTYPE); q21 = np.ascontiguousarray(q21, PROJ_DTYPE)
    d12 = np.ascontiguousarray(d12, np.uint8); d21 = np.ascontiguousarray(d21, np.uint8)
    i1 = np.ascontiguousarray(index1, np.int32); i2 = np.ascontiguousarray(index2, np.int32)
    m12 = np.full(max(kf1.n, 1), -1, np.int32)
    nf = C.c_int32()
    check(lib().orbm_search_by_sim3(C.byref(kf1.c), C.byref(kf2.c), _p(q12), _p(d12), _p(i1), len(q12), _p(q21), _p(d21),
                                    _p(i2), len(q21), device, _p(m12), C.byref(nf)))
    return nf.value, m12[:kf1.n].copy()


def search_for_triangulation(pairs, f12, epipoles, only_stereo=False, check_ori=True, device=0):
    """pairs: list of (KeyFrameData kf1, KeyFrameData kf2); f12: [npairs][9]; epipoles: [npairs][2]"""
    n = len(pairs)
    a = (KeyFrameStruct * n)(*[p[0].c for p in pairs]); b = (KeyFrameStruct * n)(*[p[1].c for p in pairs])
    F = np.ascontiguousarray(f12, np.float32).reshape(n, 9); E = np.ascontiguousarray(epipoles, np.float32).reshape(n, 2)
    tot = sum(p[0].n for p in pairs)
    m = np.full(max(tot, 1), -1, np.int32); nm = np.zeros(n, np.int32)
    check(lib().orbm_search_for_triangulation(a, b, n, _p(F), _p(E), int(only_stereo), int(check_ori), device, _p(m), _p(nm)))
    out, o = [], 0
    for p in pairs:
        out.append(m[o:o + p[0].n].copy()); o += p[0].n
    return nm, out


CLOUD_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("z", "<f4"), ("b", "u1"), ("g", "u1"), ("r", "u1"), ("a", "u1")])


def generate_point_clouds(depth, bgr, cam: "Camera", pose_inv=None, step=3, device=0, capacity=None):
    """PointCloudMapping::generatePointCloud for a batch: depth [F][H][W] float32, bgr [F][H][W][3] uint8 or None,
    pose_inv [F][4][4] float64 (T.inverse().matrix()) or None.  Returns a list of CLOUD_DTYPE arrays."""
    depth = np.ascontiguousarray(depth, np.float32)
    F, H, W = depth.shape
    if bgr is not None:
        bgr = np.ascontiguousarray(bgr, np.uint8).reshape(F, H, W, 3)
    if pose_inv is not None:
        pose_inv = np.ascontiguousarray(pose_inv, np.float64).reshape(F, 16)
    cap = capacity or (-(-W // step)) * (-(-H // step))
    pts = np.zeros((F, cap), CLOUD_DTYPE)
    counts = np.zeros(F, np.int32)
    check(lib().orbm_generate_point_clouds(_p(depth), C.c_size_t(W), C.c_size_t(W * H), _p(bgr), C.c_size_t(W * 3), C.c_size_t(W * H * 3),
                                           W, H, F, step, C.byref(cam), _p(pose_inv), device, _p(pts), cap, _p(counts)))
    return [pts[f, :counts[f]].copy() for f in range(F)]


def voxel_grid(points: np.ndarray, leaf: float, device=0, capacity=None):
    """orbm_voxel_grid: points = CLOUD_DTYPE records (host) -> one centroid per occupied voxel, ascending voxel index"""
    pts = np.ascontiguousarray(points, CLOUD_DTYPE)
    n = C.c_int32(0)
    if capacity is None:
        check(lib().orbm_voxel_grid(C.c_void_p(pts.ctypes.data), len(pts), C.c_float(leaf), device, None, 0, C.byref(n)))
        capacity = n.value
    out = np.zeros(max(capacity, 1), CLOUD_DTYPE)
    check(lib().orbm_voxel_grid(C.c_void_p(pts.ctypes.data), len(pts), C.c_float(leaf), device, C.c_void_p(out.ctypes.data), capacity, C.byref(n)))
    return out[:n.value].copy()


def statistical_outlier_removal(points: np.ndarray, mean_k: int = 50, stddev_mul: float = 1.0, device=0):
    """orbm_statistical_outlier_removal -> (kept points, mean distance per input point, [threshold, mean, stddev, valid, visited])"""
    pts = np.ascontiguousarray(points, CLOUD_DTYPE)
    out = np.zeros(max(len(pts), 1), CLOUD_DTYPE)
    md = np.zeros(max(len(pts), 1), np.float32)
    st = np.zeros(5, np.float64)
    n = C.c_int64(0)
    f = lib().orbm_statistical_outlier_removal
    f.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_double, C.c_int, C.c_void_p, C.c_int64, C.POINTER(C.c_int64), C.c_void_p, C.c_void_p]
    check(f(pts.ctypes.data_as(C.c_void_p), len(pts), int(mean_k), float(stddev_mul), device, out.ctypes.data_as(C.c_void_p), len(out),
            C.byref(n), md.ctypes.data_as(C.c_void_p), st.ctypes.data_as(C.c_void_p)))
    return out[:n.value].copy(), md[:len(pts)].copy(), st
