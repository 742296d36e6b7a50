"""Freeze the REFERENCE's outputs (oracle/_ref/libref_matchers.so = the unmodified src/ORBmatcher.cc, Frame.cc, KeyFrame.cc,
MapPoint.cc and DBoW2 compiled over oracle/ref_shim2) for every case of tests/matcher_cases.py into
tests/golden/ref_matchers.npz: full arrays for small outputs, SHA-256 for large ones.
Run where /root/reference is mounted:  python tests/golden/make_golden_matchers.py"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import matcher_cases as M  # noqa: E402
import ref2_lib as R  # noqa: E402

FULL_LIMIT = 4096          # arrays up to this many elements are stored whole


def digest(a: np.ndarray) -> np.ndarray:
    a = np.ascontiguousarray(a)
    h = hashlib.sha256()
    h.update(str(a.dtype).encode()); h.update(str(a.shape).encode()); h.update(a.tobytes())
    return np.frombuffer(h.digest(), np.uint8).copy()


def freeze(out: dict, name: str, res: dict):
    for k, v in res.items():
        v = np.asarray(v)
        if v.size <= FULL_LIMIT:
            out[f"{name}/{k}"] = v
        else:
            out[f"{name}/{k}#sha256"] = digest(v)


def main():
    assert R.available(), "build oracle/_ref/libref_matchers.so first (make -C oracle ref2)"
    out = {}
    for name, mk in M.CASES.items():
        freeze(out, name, M.run_reference(mk()))
        print("froze", name)
    np.savez_compressed(os.path.join(HERE, "ref_matchers.npz"), **out)
    print("wrote", len(out), "arrays")


if __name__ == "__main__":
    main()
