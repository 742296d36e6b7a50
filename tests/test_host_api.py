"""Host-side C++ classes (orb_slam2_map_b200/host/orbx.hpp): they compile against the C ABI (CPU
check) and, on a GPU, reproduce the oracle when driven like Frame.cc / Tracking.cc drive the
reference classes."""
import os
import struct
import subprocess

import numpy as np
import pytest

import oracle_lib as O
from orb_slam2_map_b200 import capi, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "orb_slam2_map_b200", "host")
PKG = os.path.join(ROOT, "orb_slam2_map_b200")


def _build(tmp_path):
    exe = str(tmp_path / "host_api_test")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-I", HOST, os.path.join(ROOT, "tests", "host", "host_api_test.cpp"),
                           "-o", exe, "-L", PKG, "-lorbx", "-Wl,-rpath," + PKG])
    return exe


def test_host_classes_compile_and_link(tmp_path):
    _build(tmp_path)
    shim = os.path.join(ROOT, "oracle", "ref_shim")
    subprocess.check_call(["g++", "-std=c++17", "-fsyntax-only", "-Wall", "-I", HOST, "-I", shim,
                           os.path.join(ROOT, "tests", "host", "opencv_flavour_check.cpp")])


def test_projection_flatten_rules(tmp_path):
    """ProjectionFlatten (orbx.hpp) = the reference's query expressions (src/ORBmatcher.cc:61-77, :1374-1410, :1524-1529,
    :357-377); the same rules tests/matcher_cases.py proves against the compiled reference.  Host arithmetic only."""
    exe = str(tmp_path / "flatten_check")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-I", HOST, os.path.join(ROOT, "tests", "host", "flatten_check.cpp"),
                           "-o", exe, "-L", PKG, "-lorbx", "-Wl,-rpath," + PKG])
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 0 and "flatten check ok" in r.stdout, r.stdout


def test_host_classes_fail_loudly_without_gpu(tmp_path):
    if capi.lib().orbx_device_count() > 0:
        pytest.skip("needs a machine without a CUDA device")
    exe = _build(tmp_path)
    raw = tmp_path / "in.raw"
    raw.write_bytes(struct.pack("<3i", 64, 64, 2) + bytes(64 * 64 * 2))
    r = subprocess.run([exe, str(raw), str(tmp_path / "out.bin")], capture_output=True, text=True)
    assert r.returncode == 1 and "CUDA" in r.stderr          # std::runtime_error from the constructor: no CPU fallback


@pytest.mark.gpu
def test_host_classes_match_oracle(tmp_path):
    exe = _build(tmp_path)
    w, h = 640, 480
    frames = synth.frame_sequence(2, w, h, seed=4)
    raw = tmp_path / "in.raw"
    raw.write_bytes(struct.pack("<3i", w, h, 2) + frames.tobytes())
    out = tmp_path / "out.bin"
    subprocess.check_call([exe, str(raw), str(out)])
    buf = out.read_bytes()
    pos = 0
    orc = O.Extractor()
    ref = []
    for i in range(2):
        n, = struct.unpack_from("<i", buf, pos); pos += 4
        k = np.frombuffer(buf, capi.KP_DTYPE, n, pos); pos += 28 * n
        d = np.frombuffer(buf, np.uint8, 32 * n, pos).reshape(n, 32); pos += 32 * n
        ko, do = orc.extract(frames[i])
        assert n == len(ko) and k.tobytes() == ko.tobytes() and np.array_equal(d, do)
        ref.append((ko, do))
    rows, cols = struct.unpack_from("<2i", buf, pos); pos += 8
    lvl = np.frombuffer(buf, np.uint8, rows * cols, pos).reshape(rows, cols); pos += rows * cols
    assert np.array_equal(lvl, O.border101(orc.level(3), 19))      # mvImagePyramid[3] with its 19-px border
    nm, n1 = struct.unpack_from("<2i", buf, pos); pos += 8
    m12 = np.frombuffer(buf, np.int32, n1, pos); pos += 4 * n1
    pm = np.frombuffer(buf, np.float32, 2 * n1, pos).reshape(n1, 2); pos += 8 * n1
    f1, f2 = O.FrameData(ref[0][0], ref[0][1], w, h), O.FrameData(ref[1][0], ref[1][1], w, h)
    prev = np.stack([f1.kps["x"], f1.kps["y"]], axis=1)
    n_o, m_o, p_o, _ = O.search_for_initialization(f1, f2, prev, 100, 0.9, True)
    assert nm == n_o and np.array_equal(m12, m_o) and np.array_equal(pm, p_o)
    # keyframe-rate wrappers: Fuse, ComputeDistinctiveDescriptors, generatePointCloud
    nq, = struct.unpack_from("<i", buf, pos); pos += 4
    fused = np.frombuffer(buf, np.int32, nq, pos); pos += 4 * nq
    tab = orc.tables()
    kf = O.KeyFrameData(ref[1][0], ref[1][1], w, h, None, tab["scale"], tab["sigma2"])
    kf.is2 = tab["inv_sigma2"].copy(); kf.c.inv_level_sigma2 = kf.is2.ctypes.data
    q = np.zeros(nq, O.PROJ_DTYPE)
    k0 = ref[0][0]
    q["u"], q["v"] = k0["x"], k0["y"]
    q["radius"] = np.float32(3.0) * tab["scale"][k0["octave"]]
    q["min_level"] = -1; q["max_level"] = -1; q["u_right"] = -1; q["max_right_err"] = -1; q["angle"] = k0["angle"]
    q["level_lo"] = k0["octave"] - 1; q["level_hi"] = k0["octave"]
    bo, _ = O.fuse(0, kf, q, ref[0][1])
    assert np.array_equal(fused, bo) and (fused >= 0).sum() > 50
    npts, = struct.unpack_from("<i", buf, pos); pos += 4
    dd = np.frombuffer(buf, np.int32, npts, pos); pos += 4 * npts
    offs = np.arange(0, len(k0) - 4, 5, dtype=np.int32)[:101]
    io, _ = O.distinctive_descriptors(ref[0][1], offs)
    assert npts == len(offs) - 1 and np.array_equal(dd, io)
    nc, = struct.unpack_from("<i", buf, pos); pos += 4
    cloud = np.frombuffer(buf, O.CLOUD_DTYPE, nc, pos); pos += 16 * nc
    Tinv = np.array([0.36, 0.48, -0.8, 0.1, -0.8, 0.6, 0, -0.2, 0.48, 0.64, 0.6, 0.3, 0, 0, 0, 1], np.float64)
    exp = O.generate_point_cloud(frames[0].astype(np.float32) / np.float32(32.0), None, np.float32(517.3), np.float32(516.5),
                                 np.float32(318.6), np.float32(255.3), pose_inv=Tinv)
    assert nc == len(exp) and cloud.tobytes() == exp.tobytes()
    nv, = struct.unpack_from("<i", buf, pos); pos += 4
    vox = np.frombuffer(buf, O.CLOUD_DTYPE, nv, pos); pos += 16 * nv
    vo = O.voxel_grid(exp, 0.04)
    assert nv == len(vo) and vox.tobytes() == vo.tobytes() and nv > 1000
    nk, = struct.unpack_from("<i", buf, pos); pos += 4
    kept = np.frombuffer(buf, O.CLOUD_DTYPE, nk, pos); pos += 16 * nk
    ko, _, _ = O.statistical_outlier_removal(vo, 50, 1.0)
    assert nk == len(ko) and kept.tobytes() == ko.tobytes() and 0 < nk < nv
    assert pos == len(buf)
