"""PointCloudMapping's voxel-grid down-sampling (pcl::VoxelGrid, src/PointCloudMap.cc:41, 240-241, 259-260, 277-280).

PCL is not vendored under /root/reference: the oracle (oracle/frame_oracle.cpp: oracle_voxel_grid) restates its published
algorithm — parity unpinned.  CPU test: the oracle against an independent numpy statement (occupied voxels, their order, integer
colour means exact; float centroids within 1e-5 relative — numpy sums pairwise).  GPU test: orbm_voxel_grid == oracle bit for bit
(both add a voxel's points in input order), on random clouds, a back-projected depth map, duplicated points, non-finite points,
one-voxel and one-point clouds, and a grid too fine for PCL's int voxel index."""
import numpy as np
import pytest

import oracle_lib as O


def _cloud(n, seed, spread=1.0):
    r = np.random.default_rng(seed)
    p = np.zeros(n, O.CLOUD_DTYPE)
    p["x"] = r.normal(0, spread, n); p["y"] = r.normal(0, spread, n); p["z"] = r.uniform(0.4, 4.0, n)
    for c in "bgr":
        p[c] = r.integers(0, 256, n)
    p["a"] = 255
    return p


def _numpy_voxel(p, leaf):
    ok = np.isfinite(p["x"]) & np.isfinite(p["y"]) & np.isfinite(p["z"])
    q = p[ok]
    inv = np.float32(1.0) / np.float32(leaf)
    xyz = np.stack([q["x"], q["y"], q["z"]], 1)
    minb = np.floor(xyz.min(0) * inv).astype(np.int64)
    div = np.floor(xyz.max(0) * inv).astype(np.int64) - minb + 1
    ijk = (np.floor(xyz * inv) - minb.astype(np.float32)).astype(np.int64)
    idx = ijk[:, 0] + ijk[:, 1] * div[0] + ijk[:, 2] * div[0] * div[1]
    keys, invmap, cnt = np.unique(idx, return_inverse=True, return_counts=True)
    out = np.zeros(len(keys), O.CLOUD_DTYPE)
    for c in "xyz":
        out[c] = (np.bincount(invmap, q[c].astype(np.float64)) / cnt).astype(np.float32)
    for c in "bgr":
        s = np.bincount(invmap, q[c].astype(np.int64)).astype(np.int64)
        out[c] = (s.astype(np.float32) / cnt.astype(np.float32)).astype(np.int32).astype(np.uint8)
    out["a"] = 255
    return out


@pytest.mark.parametrize("n,leaf,seed", [(20000, 0.05, 1), (20000, 0.5, 2), (300, 0.01, 3), (5000, 3.0, 4)])
def test_oracle_voxel_grid_vs_numpy(n, leaf, seed):
    p = _cloud(n, seed)
    p["x"][::97] = np.nan
    v, w = O.voxel_grid(p, leaf), _numpy_voxel(p, leaf)
    assert len(v) == len(w)
    for c in "bgra":
        assert np.array_equal(v[c], w[c]), c
    for c in "xyz":
        assert np.allclose(v[c], w[c], rtol=1e-5, atol=1e-6), c


def test_oracle_voxel_grid_edges():
    assert len(O.voxel_grid(np.zeros(0, O.CLOUD_DTYPE), 0.1)) == 0
    one = _cloud(1, 5)
    assert O.voxel_grid(one, 0.1).tobytes() == one.tobytes()
    same = np.repeat(_cloud(1, 6), 50)
    v = O.voxel_grid(same, 0.2)
    assert len(v) == 1 and v["b"][0] == same["b"][0]
    nan = _cloud(10, 7); nan["z"] = np.inf
    assert len(O.voxel_grid(nan, 0.1)) == 0


@pytest.mark.gpu
@pytest.mark.parametrize("n,leaf,seed", [(200000, 0.05, 1), (200000, 0.4, 2), (1000, 0.01, 3), (50000, 5.0, 4), (1, 0.1, 5)])
def test_voxel_grid_gpu_equals_oracle(n, leaf, seed):
    from orb_slam2_map_b200 import capi
    p = _cloud(n, seed)
    if n > 100:
        p["y"][::53] = np.nan
        p[10:40] = p[5]                                       # duplicated points in one voxel
    vo = O.voxel_grid(p, leaf)
    vg = capi.voxel_grid(p, leaf)
    assert len(vg) == len(vo)
    assert vg.tobytes() == vo.tobytes()


@pytest.mark.gpu
def test_voxel_grid_gpu_on_backprojected_keyframes_and_limits():
    from orb_slam2_map_b200 import capi
    r = np.random.default_rng(9)
    depth = (1.0 + 2.0 * r.random((480, 640))).astype(np.float32)
    depth[r.random(depth.shape) < 0.1] = 0
    bgr = r.integers(0, 256, (480, 640, 3), dtype=np.uint8)
    cloud = O.generate_point_cloud(depth, bgr, 517.3, 516.5, 318.6, 255.3, None, 3)
    both = np.concatenate([cloud, cloud[::2]])                # "globalMap += keyframe cloud", then filter (PointCloudMap.cc:236-241)
    for leaf in (0.01, 0.04):
        vo, vg = O.voxel_grid(both, leaf), capi.voxel_grid(both, leaf)
        assert len(vo) > 1000 and vg.tobytes() == vo.tobytes()
    # capacity too small and PCL's index-overflow refusal
    with pytest.raises(capi.OrbxError):
        capi.voxel_grid(both, 0.04, capacity=10)
    far = _cloud(100, 1, spread=500.0)
    with pytest.raises(capi.OrbxError):
        capi.voxel_grid(far, 0.001)
    assert len(capi.voxel_grid(np.zeros(0, O.CLOUD_DTYPE), 0.1)) == 0
