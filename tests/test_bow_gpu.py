"""GPU parity of the bag-of-words path (orbx_bow.cu through the C ABI) against oracle/bow_oracle.cpp: word / weight /
node ids of DBoW2's per-feature transform and both SearchByBoW variants, bit-exact, batched."""
import numpy as np
import pytest

import oracle_lib as O
from orb_slam2_map_b200 import capi
from test_oracle_bow import make_pair

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("k,depth", [(10, 3), (5, 4), (33, 2)])
def test_bow_transform(k, depth):
    voc = O.make_vocabulary(k=k, depth=depth, seed=k)
    dv = capi.Vocabulary(voc)
    rng = np.random.default_rng(1)
    feats = rng.integers(0, 256, (700, 32), dtype=np.uint8)
    leaves = np.flatnonzero(voc["child_count"] == 0)
    feats[:400] = voc["desc"][rng.choice(leaves, 400)]
    feats[400:500] = voc["desc"][rng.integers(1, voc["nnodes"], 100)]      # inner-node descriptors: distance ties between children
    for levelsup in (0, 1, 2, 4, 9):
        wg, vg, ng = dv.transform(feats, levelsup)
        wo, vo, no = O.bow_transform(voc, feats, levelsup)
        assert np.array_equal(wg, wo) and np.array_equal(vg, vo) and np.array_equal(ng, no), levelsup
    dv.close()


@pytest.mark.parametrize("variant", [0, 1])
def test_search_by_bow_batch(variant):
    voc = O.make_vocabulary(k=6, depth=3, seed=3)
    pairs = []
    for seed, n, lu in [(0, 220, 1), (1, 1000, 2), (2, 37, 1), (3, 1500, 3), (4, 640, 2), (5, 4500, 2)]:   # 4500: beyond the register sort
        pairs.append(make_pair(voc, n=n, seed=seed, levelsup=lu))
    # an empty frame and a pair without common nodes
    e = (np.zeros((0, 32), np.uint8), np.zeros(0, np.float32), np.zeros(0, np.uint32), np.zeros(0, np.uint8))
    pairs.append((e, pairs[0][1]))
    a, b = pairs[0]
    pairs.append((a, (b[0], b[1], b[2] + 100000, b[3])))
    for ratio, ori in [(0.7, True), (0.9, True), (0.9, False)]:
        got = capi.search_by_bow(variant, pairs, nnratio=ratio, check_ori=ori)
        total = 0
        for (A, B), (n, mA, mB) in zip(pairs, got):
            n_o, mA_o, mB_o = O.search_by_bow(variant, *A, *B, nnratio=ratio, check_ori=ori)
            assert n == n_o and np.array_equal(mA, mA_o) and np.array_equal(mB, mB_o)
            total += n
        assert total > 30


def test_distinctive_descriptors():
    rng = np.random.default_rng(8)
    sizes = [1, 2, 3, 7, 0, 20, 64, 5, 130, 300, 33, 2500]
    offsets = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    base = rng.integers(0, 256, (len(sizes), 32), dtype=np.uint8)
    desc = np.concatenate([np.repeat(base[i][None], n, 0) for i, n in enumerate(sizes)])
    desc ^= rng.integers(0, 256, desc.shape, dtype=np.uint8) & rng.integers(0, 256, desc.shape, dtype=np.uint8) & rng.integers(0, 256, desc.shape, dtype=np.uint8)
    ig, mg = capi.distinctive_descriptors(desc, offsets)
    io, mo = O.distinctive_descriptors(desc, offsets)
    assert np.array_equal(ig, io) and np.array_equal(mg, mo)
