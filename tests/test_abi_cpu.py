"""CPU tests (no GPU): the C-ABI library loads, exports every symbol include/orbx.h declares, has
the documented struct layouts, and fails LOUDLY (no CPU fallback) when no CUDA device is present."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from orb_slam2_map_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "orbx.h")


def _declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = re.findall(r"\b(orb[xm]_[a-z0-9_]+)\s*\(", src)
    return sorted(set(names))


def test_header_symbols_are_exported():
    names = _declared_functions()
    assert len(names) >= 20
    L = capi.lib()
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/orbx.h but not exported by liborbx.so"
    # and the ctypes plumbing knows all of them
    assert set(names) <= set(capi.EXPORTS), set(names) - set(capi.EXPORTS)


def test_dynamic_symbol_table_matches_header():
    out = subprocess.check_output(["nm", "-D", "--defined-only", capi.LIB_PATH], text=True)
    exported = {ln.split()[-1] for ln in out.splitlines() if " T " in ln}
    for n in _declared_functions():
        assert n in exported, n
    # the product must not link or embed the oracle
    assert not any(s.startswith("oracle_") for s in exported)
    deps = subprocess.check_output(["ldd", capi.LIB_PATH], text=True)
    assert "oracle" not in deps


def test_struct_layouts():
    assert capi.KP_DTYPE.itemsize == 28                       # cv::KeyPoint
    assert capi.PROJ_DTYPE.itemsize == 40
    assert C.sizeof(capi.Params) == 20
    assert C.sizeof(capi.FrameStruct) == 64
    assert C.sizeof(capi.StereoInput) == 120


def test_header_is_plain_c_and_layouts_agree(tmp_path):
    """include/orbx.h compiles as C (no C++/torch types) and gcc's layouts equal the ctypes ones."""
    src = tmp_path / "t.c"
    src.write_text('#include <stdio.h>\n#include "orbx.h"\nint main(void){printf("%zu %zu %zu %zu %zu\\n",'
                   'sizeof(orbx_keypoint),sizeof(orbx_params),sizeof(orbm_frame),sizeof(orbm_proj_query),'
                   'sizeof(orbm_stereo_input));return 0;}\n')
    exe = tmp_path / "t"
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    sizes = [int(x) for x in subprocess.check_output([str(exe)], text=True).split()]
    assert sizes == [capi.KP_DTYPE.itemsize, C.sizeof(capi.Params), C.sizeof(capi.FrameStruct),
                     capi.PROJ_DTYPE.itemsize, C.sizeof(capi.StereoInput)]


def _no_gpu():
    return capi.lib().orbx_device_count() <= 0


@pytest.mark.skipif(not _no_gpu(), reason="only meaningful on a machine without a CUDA device")
def test_no_cpu_fallback_without_device():
    L = capi.lib()
    with pytest.raises(capi.OrbxError) as e:
        capi.Extractor()
    assert "-2" in str(e.value) or "CUDA" in str(e.value)
    a = np.zeros((4, 32), np.uint8)
    with pytest.raises(capi.OrbxError):
        capi.descriptor_distance(a, a)
    k = np.zeros(3, capi.KP_DTYPE)
    f = capi.FrameData(k, np.zeros((3, 32), np.uint8), 640, 480)
    with pytest.raises(capi.OrbxError):
        capi.search_for_initialization(f, f, np.zeros((3, 2), np.float32))
    assert L.orbx_last_error()          # message available to the caller


def test_argument_validation_needs_no_device():
    L = capi.lib()
    h = C.c_void_p()
    bad = capi.Params(1000, 0.9, 8, 20, 7)                   # scale factor <= 1
    assert L.orbx_create(C.byref(bad), 640, 480, 1, 1, 0, None, C.byref(h)) == -1
    assert b"scale" in L.orbx_last_error()
    assert L.orbx_create(None, 640, 480, 1, 1, 0, None, C.byref(h)) == -1
    assert L.orbx_sync(None) == -1
    assert L.orbx_destroy(None) == 0
    # empty image: silent return, outputs untouched (reference src/ORBextractor.cc:1046-1047)
    n = C.c_int(-5)
    assert L.orbx_extract(None, None, 0, 0, 0, None, None, 0, C.byref(n)) in (0, -1)
    assert n.value == -5


def test_package_never_imports_oracle():
    """The product package must not reference oracle/ (CPU fallback would void parity claims)."""
    pkg = os.path.join(ROOT, "orb_slam2_map_b200")
    for dp, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".h", ".cpp", ".hpp", ".cuh")) or fn == "Makefile":
                txt = open(os.path.join(dp, fn), errors="ignore").read()
                assert "oracle_lib" not in txt and "orb_oracle" not in txt and "liborb_oracle" not in txt, os.path.join(dp, fn)
