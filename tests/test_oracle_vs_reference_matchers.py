"""Pins the matcher half of the oracle to the reference's OWN code.

oracle/_ref/libref_matchers.so is the unmodified /root/reference/src/{ORBmatcher,Frame,KeyFrame,MapPoint,Map,
KeyFrameDatabase,ORBextractor}.cc + Thirdparty/DBoW2 compiled over oracle/ref_shim2 (typed cv stand-in, stub pcl / boost /
Eigen / g2o headers) and driven from flat arrays by oracle/ref_shim2/ref_driver2.cpp.  For every case of
tests/matcher_cases.py:

  test_oracle_equals_reference   oracle == reference, live (runs where the library could be built: /root/reference
                                 mounted, or a prebuilt copy travelled with gpurun)
  test_oracle_equals_golden      oracle == tests/golden/ref_matchers.npz, the reference's outputs frozen by
                                 tests/golden/make_golden_matchers.py (runs everywhere)

Bit-exact: indices, counts, Hamming distances, and the float outputs of ComputeStereoMatches."""
import os

import numpy as np
import pytest

import matcher_cases as M
import ref2_lib as R
from golden.make_golden_matchers import FULL_LIMIT, digest

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_matchers.npz")


@pytest.mark.skipif(not R.available(), reason="oracle/_ref/libref_matchers.so not built (needs /root/reference)")
@pytest.mark.parametrize("name", list(M.CASES))
def test_oracle_equals_reference(name):
    c = M.CASES[name]()
    ref, orc = M.run_reference(c), M.run_oracle(c)
    assert set(ref) == set(orc)
    for k in ref:
        assert np.array_equal(ref[k], orc[k]), f"{name}: {k} differs from the reference"
    if "n" in ref:
        assert int(ref["n"]) > 10, "case too easy: the reference found almost nothing"


@pytest.mark.parametrize("name", list(M.CASES))
def test_oracle_equals_golden(name):
    g = np.load(GOLDEN)
    orc = M.run_oracle(M.CASES[name]())
    seen = 0
    for k, v in orc.items():
        v = np.asarray(v)
        if v.size <= FULL_LIMIT:
            assert np.array_equal(g[f"{name}/{k}"], v), f"{name}: {k} differs from the frozen reference output"
        else:
            assert np.array_equal(g[f"{name}/{k}#sha256"], digest(v)), f"{name}: {k} differs from the frozen reference output"
        seen += 1
    assert seen == sum(1 for key in g.files if key.startswith(name + "/"))
