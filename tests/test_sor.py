"""PointCloudMapping's statistical outlier removal (pcl::StatisticalOutlierRemoval, src/PointCloudMap.cc:46-47, 284-285).

PCL is not vendored under /root/reference: the oracle (oracle/frame_oracle.cpp: oracle_statistical_outlier_removal) restates its
published algorithm — parity unpinned.  CPU test: the oracle against an independent numpy statement (full float32 distance matrix
in FLANN's accumulation order, row sort, sequential double sums): mean distances, threshold and kept set identical.  GPU test:
orbm_statistical_outlier_removal == oracle: mean distances bit for bit (integer-like work: the k smallest float distances are a
set, their square roots are summed in ascending order in double), kept points identical; the threshold is compared within 1e-12
relative because the device adds the distances in a fixed tree order where PCL's loop is serial (no test point lies that close to
the threshold, which the test checks)."""
import numpy as np
import pytest

import oracle_lib as O


def _cloud(n, seed, outliers=0.03, coherent=True):
    """points on a few surfaces + far outliers; coherent = in voxel-index order (what the voxel filter emits)"""
    r = np.random.default_rng(seed)
    p = np.zeros(n, O.CLOUD_DTYPE)
    k = r.integers(0, 3, n)
    u, v = r.uniform(-1.5, 1.5, n), r.uniform(-1.0, 1.0, n)
    p["x"] = np.where(k == 0, u, np.where(k == 1, -1.5, u)); p["y"] = np.where(k == 2, -1.0, v)
    p["z"] = np.where(k == 0, 3.0, np.where(k == 1, 2.0 + u, 2.0 + v)) + r.normal(0, 0.004, n)
    far = r.random(n) < outliers
    for c in "xyz":
        p[c][far] = r.uniform(-4, 6, int(far.sum()))
    for c in "bgr":
        p[c] = r.integers(0, 256, n)
    p["a"] = 255
    if coherent:
        key = (np.floor(p["z"] / 0.25) * 4096 + np.floor(p["y"] / 0.25 + 64)) * 4096 + np.floor(p["x"] / 0.25 + 64)
        p = p[np.argsort(key, kind="stable")]
    return p


def _numpy_sor(p, mean_k, mul):
    ok = np.isfinite(p["x"]) & np.isfinite(p["y"]) & np.isfinite(p["z"])
    n = len(p)
    dist = np.zeros(n, np.float32)
    idx = np.nonzero(ok)[0]
    if len(idx) >= mean_k + 1:
        x, y, z = p["x"][idx], p["y"][idx], p["z"][idx]
        for a, i in enumerate(idx):
            dx, dy, dz = x[a] - x, y[a] - y, z[a] - z
            d2 = np.sort((dx * dx + dy * dy) + dz * dz)[:mean_k + 1]          # float32 throughout, left to right
            dist[i] = np.float32(np.cumsum(np.sqrt(d2[1:].astype(np.float64)))[-1] / mean_k)
    valid = float(len(idx)) if len(idx) >= mean_k + 1 else 0.0
    s = np.cumsum(dist.astype(np.float64))[-1] if n else 0.0                   # cumsum adds serially, like PCL's loop
    q = np.cumsum((dist * dist).astype(np.float64))[-1] if n else 0.0
    with np.errstate(all="ignore"):
        mean = np.float64(s) / valid
        var = (q - s * s / valid) / (valid - 1.0)
        thr = mean + mul * np.sqrt(var)
    keep = ~(dist.astype(np.float64) > thr)
    return p[keep], dist, thr


@pytest.mark.parametrize("n,mean_k,mul,seed", [(1500, 50, 1.0, 1), (700, 8, 2.0, 2), (52, 50, 1.0, 3), (300, 1, 0.5, 4)])
def test_oracle_sor_vs_numpy(n, mean_k, mul, seed):
    p = _cloud(n, seed)
    p["x"][::131] = np.nan
    p[5] = p[6]                                                   # a duplicate: a neighbour at distance 0
    kept, md, st = O.statistical_outlier_removal(p, mean_k, mul)
    kept_n, md_n, thr_n = _numpy_sor(p, mean_k, mul)
    assert np.array_equal(md, md_n)
    assert st[0] == thr_n and kept.tobytes() == kept_n.tobytes()
    assert 0 < len(kept) < n


def test_oracle_sor_edges():
    assert len(O.statistical_outlier_removal(np.zeros(0, O.CLOUD_DTYPE))[0]) == 0
    few = _cloud(40, 7)                                           # fewer than mean_k + 1 points: the search fails, everything stays
    kept, md, st = O.statistical_outlier_removal(few, 50, 1.0)
    assert kept.tobytes() == few.tobytes() and not md.any() and st[3] == 0 and np.isnan(st[0])


def _gpu_equals_oracle(p, mean_k, mul):
    from orb_slam2_map_b200 import capi
    kept_o, md_o, st_o = O.statistical_outlier_removal(p, mean_k, mul)
    kept_g, md_g, st_g = capi.statistical_outlier_removal(p, mean_k, mul)
    assert np.array_equal(md_g, md_o)
    if np.isnan(st_o[0]):
        assert np.isnan(st_g[0])
    else:
        assert abs(st_g[0] - st_o[0]) <= 1e-12 * abs(st_o[0])
        assert np.abs(md_o.astype(np.float64) - st_o[0]).min() > 1e-9       # nobody sits on the threshold
    assert st_g[3] == st_o[3] and kept_g.tobytes() == kept_o.tobytes()
    return st_g


@pytest.mark.gpu
@pytest.mark.parametrize("n,mean_k,mul,seed,coherent", [(6000, 50, 1.0, 1, True), (6000, 50, 1.0, 2, False), (900, 8, 2.0, 3, True),
                                                       (257, 127, 1.0, 4, True), (52, 50, 1.0, 5, False), (51, 50, 1.0, 6, True)])
def test_sor_gpu_equals_oracle(n, mean_k, mul, seed, coherent):
    p = _cloud(n, seed, coherent=coherent)
    p["y"][::113] = np.inf
    p[7] = p[8]
    _gpu_equals_oracle(p, mean_k, mul)


@pytest.mark.gpu
def test_sor_gpu_edges_and_tile_skipping():
    from orb_slam2_map_b200 import capi
    assert len(capi.statistical_outlier_removal(np.zeros(0, O.CLOUD_DTYPE))[0]) == 0
    few = _cloud(40, 7)
    kept, md, st = capi.statistical_outlier_removal(few, 50, 1.0)
    assert kept.tobytes() == few.tobytes() and not md.any() and np.isnan(st[0])
    with pytest.raises(capi.OrbxError):
        capi.statistical_outlier_removal(few, 128, 1.0)
    # a coherent cloud (voxel-index order, as the voxel filter emits it) lets the search skip most tile pairs; the result is the same
    p = _cloud(20000, 9, coherent=True)
    st = _gpu_equals_oracle(p, 50, 1.0)
    assert st[4] < 0.5, st[4]
