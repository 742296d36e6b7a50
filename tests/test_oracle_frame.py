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
