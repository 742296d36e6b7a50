"""The CUDA matcher path against the REFERENCE's own frozen outputs, with nothing in between: every case of
tests/matcher_cases.py goes through the C ABI (orbm_* calls, flattened exactly as the oracle's inputs are) and is compared with
tests/golden/ref_matchers.npz, the outputs of the unmodified src/ORBmatcher.cc / Frame.cc / KeyFrame.cc / MapPoint.cc / DBoW2
compiled here (oracle/_ref/libref_matchers.so, tests/golden/make_golden_matchers.py).  Bit-exact: indices, counts, distances,
uRight / depth floats, bag-of-words weights.  Frame::GetFeaturesInArea has no entry point of its own (every search contains it)."""
import os

import numpy as np
import pytest

import matcher_cases as M
import oracle_lib as O
from golden.make_golden_matchers import FULL_LIMIT, digest
from orb_slam2_map_b200 import capi

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_matchers.npz")


class Gpu:
    """the functions run_oracle() calls, over liborbx.so"""
    FrameData = capi.FrameData
    KeyFrameData = capi.KeyFrameData
    compute_stereo_matches = staticmethod(capi.compute_stereo_matches)
    search_by_projection = staticmethod(capi.search_by_projection)
    fuse = staticmethod(capi.fuse)
    search_by_sim3 = staticmethod(capi.search_by_sim3)
    distinctive_descriptors = staticmethod(capi.distinctive_descriptors)

    @staticmethod
    def descriptor_distance(a, b):
        return int(capi.descriptor_distance(np.asarray(a)[None], np.asarray(b)[None])[0])

    @staticmethod
    def search_for_initialization(f1, f2, prev, window, nnratio, ori):
        n, m12, pm = capi.search_for_initialization(f1, f2, prev, window, nnratio, ori)
        return n, m12, pm, None

    @staticmethod
    def bow_transform(voc, feats, levelsup):
        dv = capi.Vocabulary(voc)
        try:
            return dv.transform(feats, levelsup)
        finally:
            dv.close()

    @staticmethod
    def search_by_bow(variant, dA, angA, nodeA, flA, dB, angB, nodeB, flB, nnratio, ori):
        (n, ma, mb), = capi.search_by_bow(variant, [((dA, angA, nodeA, flA), (dB, angB, nodeB, flB))], nnratio=nnratio, check_ori=ori)
        return n, ma, mb

    @staticmethod
    def search_for_triangulation(kf1, kf2, f12, epipole, only_stereo, ori):
        nm, ms = capi.search_for_triangulation([(kf1, kf2)], [f12], [epipole], only_stereo, ori)
        return int(nm[0]), ms[0]


@pytest.mark.parametrize("name", [n for n in M.CASES if not n.startswith("area")])
def test_cuda_path_equals_frozen_reference_output(name):
    g = np.load(GOLDEN)
    out = M.run_oracle(M.CASES[name](), B=Gpu)
    seen = 0
    for k, v in out.items():
        v = np.asarray(v)
        if v.size <= FULL_LIMIT:
            assert np.array_equal(g[f"{name}/{k}"], v), f"{name}: {k} differs from the reference's frozen output"
        else:
            assert np.array_equal(g[f"{name}/{k}#sha256"], digest(v)), f"{name}: {k} differs from the reference's frozen output"
        seen += 1
    assert seen == sum(1 for key in g.files if key.startswith(name + "/"))


# ---- the extractor against tests/golden/ref_orbextractor.npz (the unmodified src/ORBextractor.cc, oracle/_ref) ----------------
import hashlib
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import make_golden as G  # noqa: E402  (input generators + case tables only)

REF = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_orbextractor.npz"))


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_cuda_extractor_equals_frozen_reference_output_full():
    for i, (w, h, nf, sc, nl, it, mt, seed) in enumerate(G.EXTRACT_FULL):
        ex = capi.Extractor(nf, sc, nl, it, mt, max_width=w, max_height=h)
        k, d = ex.extract(G.extract_input(w, h, seed))
        assert np.array_equal(k.view(np.uint8).reshape(len(k), 28), REF[f"full_kps_{i}"]), i
        assert np.array_equal(d, REF[f"full_desc_{i}"]), i
        assert np.array_equal(ex.level(0, min(3, nl - 1), border=19), REF[f"full_level3_border_{i}"]), i      # mvImagePyramid with its border
        ex.close()


def test_cuda_extractor_equals_frozen_reference_output_hashes():
    for row, (w, h, nf, sc, nl, it, mt, seed) in zip(REF["hash"], G.EXTRACT_HASH):
        ex = capi.Extractor(nf, sc, nl, it, mt, max_width=w, max_height=h)
        k, d = ex.extract(G.extract_input(w, h, seed))
        assert len(k) == int(row[0])
        assert _sha(k) == str(row[1]) and _sha(d) == str(row[2])
        assert _sha(np.concatenate([ex.level(0, l).ravel() for l in range(nl)])) == str(row[3])
        ex.close()
    t = capi.Extractor().tables()
    for key in ("scale", "inv_scale", "sigma2", "inv_sigma2"):
        assert np.array_equal(t[key], REF["table_" + key]), key
