"""The oracle's RGB-D frame-front restatements (oracle/frame_oracle.cpp) against golden vectors produced by the
cv2 4.13.0 wheel (tests/golden/frame_front.npz, generator make_golden_frame.py), live against cv2 on fresh seeds
where it is importable, and against independent numpy statements of Frame.cc's loops."""
import os

import numpy as np
import pytest

import oracle_lib as O

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "frame_front.npz"))


def test_cvt_gray_golden():
    assert np.array_equal(O.cvt_gray(G["rgb"], True), G["gray_rgb"])
    assert np.array_equal(O.cvt_gray(G["rgb"], False), G["gray_bgr"])
    assert np.array_equal(O.cvt_gray(G["rgba"], True), G["gray_rgba"])
    assert np.array_equal(O.cvt_gray(G["rgba"], False), G["gray_bgra"])


def test_undistort_points_golden():
    for i in range(2):
        got = O.undistort_points(G["pts"], G["cams"][i], G["dists"][i])
        assert np.array_equal(got, G[f"undist{i}"]), np.abs(got - G[f"undist{i}"]).max()


def test_live_cv2_if_available():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(77)
    img = rng.integers(0, 256, (37, 53, 3), dtype=np.uint8)
    assert np.array_equal(O.cvt_gray(img, True), cv2.cvtColor(img, cv2.COLOR_RGB2GRAY))
    assert np.array_equal(O.cvt_gray(img, False), cv2.cvtColor(img, cv2.COLOR_BGR2GRAY))
    cam = np.array([535.4, 539.2, 320.1, 247.6], np.float32)
    dist = np.array([0.231222, -0.784899, -0.003257, -0.000105, 0.917205], np.float32)     # TUM3-like / TUM2.yaml
    K = np.array([[cam[0], 0, cam[2]], [0, cam[1], cam[3]], [0, 0, 1]], np.float32)
    pts = (rng.random((500, 2)) * [640, 480]).astype(np.float32)
    ref = cv2.undistortPoints(pts.reshape(-1, 1, 2), K, dist, None, K).reshape(-1, 2)
    assert np.array_equal(O.undistort_points(pts, cam, dist), ref)


def test_image_bounds():
    cam, dist = G["cams"][0], G["dists"][0]
    b = O.image_bounds(640, 480, cam, dist)
    u = G["undist0"][:4]            # corners (0,0) (640,0) (0,480) (640,480), Frame.cc:440-454
    assert b[0] == min(u[0, 0], u[2, 0]) and b[1] == max(u[1, 0], u[3, 0])
    assert b[2] == min(u[0, 1], u[1, 1]) and b[3] == max(u[2, 1], u[3, 1])
    assert np.array_equal(O.image_bounds(640, 480, cam, np.zeros(5, np.float32)), np.array([0, 640, 0, 480], np.float32))


def test_depth_and_stereo_from_rgbd():
    rng = np.random.default_rng(3)
    d16 = rng.integers(0, 40000, (48, 64), dtype=np.uint16)
    d16[rng.random(d16.shape) < 0.2] = 0
    f = np.float32(1.0) / np.float32(5000.0)        # mDepthMapFactor = 1.0f / DepthMapFactor, Tracking.cc:147-151
    depth = O.depth_to_float(d16, f)
    assert np.array_equal(depth, d16.astype(np.float32) * f)
    keys = (rng.random((300, 2)) * [63, 47]).astype(np.float32)
    keys_un = keys + rng.normal(0, 0.5, keys.shape).astype(np.float32)
    bf = np.float32(40.0)
    ur, z = O.stereo_from_rgbd(keys, keys_un, depth, bf)
    for i in range(len(keys)):                      # Frame.cc:646-661
        d = depth[int(keys[i, 1]), int(keys[i, 0])]
        if d > 0:
            assert z[i] == d and ur[i] == np.float32(keys_un[i, 0] - np.float32(bf / d))
        else:
            assert z[i] == -1 and ur[i] == -1


def make_cloud_case(seed, w=640, h=480):
    rng = np.random.default_rng(seed)
    depth = rng.uniform(0.3, 9.0, (h, w)).astype(np.float32)
    depth[rng.random((h, w)) < 0.25] = 0.0                      # holes
    depth[rng.random((h, w)) < 0.05] = np.float32(12.5)         # beyond 10 m
    depth[0, 0], depth[0, 3], depth[3, 0], depth[3, 3] = np.float32(0.01), np.float32(0.0099999), np.float32(10.0), np.float32(10.000001)
    bgr = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    a = rng.normal(size=(3, 3)); q, _ = np.linalg.qr(a)
    T = np.eye(4); T[:3, :3] = q * np.sign(np.linalg.det(q)); T[:3, 3] = rng.normal(0, 2, 3)
    return depth, bgr, np.linalg.inv(T)


def cloud_numpy(depth, bgr, fx, fy, cx, cy, pose_inv, step=3):
    """src/PointCloudMap.cc:78-110 vectorised in float32 / float64 numpy"""
    f32 = np.float32
    d = depth[::step, ::step]
    m, n = np.meshgrid(np.arange(0, depth.shape[0], step), np.arange(0, depth.shape[1], step), indexing="ij")
    keep = ~((d.astype(np.float64) < 0.01) | (d > f32(10)))
    z = d[keep]
    x = ((n[keep].astype(f32) - f32(cx)) * z / f32(fx)).astype(f32)
    y = ((m[keep].astype(f32) - f32(cy)) * z / f32(fy)).astype(f32)
    out = np.zeros(len(z), O.CLOUD_DTYPE)
    if pose_inv is not None:
        P = np.stack([x, y, z]).astype(np.float64)
        T = pose_inv
        for k, name in enumerate("xyz"):
            out[name] = (((T[k, 0] * P[0] + T[k, 1] * P[1]) + T[k, 2] * P[2]) + T[k, 3]).astype(f32)
    else:
        out["x"], out["y"], out["z"] = x, y, z
    out["a"] = 255
    if bgr is not None:
        c = bgr[::step, ::step][keep]
        out["b"], out["g"], out["r"] = c[:, 0], c[:, 1], c[:, 2]
    return out


def test_generate_point_cloud_vs_numpy():
    depth, bgr, Tinv = make_cloud_case(3)
    cam = (517.3, 516.5, 318.6, 255.3)
    for pose, col in [(Tinv, bgr), (None, bgr), (Tinv, None)]:
        got = O.generate_point_cloud(depth, col, *cam, pose_inv=pose)
        exp = cloud_numpy(depth, col, *cam, pose)
        assert len(got) == len(exp) and got.tobytes() == exp.tobytes()
    assert 0 < len(got) < 214 * 160
    pts = O.generate_point_cloud(depth, bgr, *cam)
    # float(0.01f) < 0.01 in double, so a depth of exactly 0.01f is dropped; 10.0f is kept
    assert pts["z"][0] != np.float32(0.01) and pts["z"].max() == np.float32(10.0)


def test_write_keyframe_features_layout():
    rng = np.random.default_rng(4)
    n = 57
    k = np.zeros(n, O.KP_DTYPE)
    for name in ("x", "y", "size", "angle", "response"):
        k[name] = rng.uniform(0, 500, n).astype(np.float32)
    k["octave"] = rng.integers(0, 8, n); k["class_id"] = -1
    d = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    mp = rng.integers(0, 1 << 40, n).astype(np.uint64); mp[::3] = np.uint64(0xFFFFFFFFFFFFFFFF)
    rec = O.write_keyframe_features(k, d, mp)
    dt = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"),
                   ("desc", "u1", 32), ("mp", "<u8")])
    assert dt.itemsize == 64
    r = rec.reshape(-1).view(dt)
    for name in ("x", "y", "size", "angle", "response", "octave"):
        assert np.array_equal(r[name], k[name])
    assert np.array_equal(r["desc"], d) and np.array_equal(r["mp"], mp)
    assert (O.write_keyframe_features(k, d, None).reshape(-1).view(dt)["mp"] == np.uint64(0xFFFFFFFFFFFFFFFF)).all()
