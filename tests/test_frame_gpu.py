"""GPU parity of the RGB-D frame front (orbx_frame.cu, through the C ABI) against oracle/frame_oracle.cpp, which is
pinned to cv2 4.13 by tests/golden/frame_front.npz: colour conversion and undistortion bit-exact, the fused batch
(colour -> gray -> ORB -> mvKeysUn / mvuRight / mvDepth -> grid -> matcher on undistorted keypoints) bit-exact."""
import os

import numpy as np
import pytest

import oracle_lib as O
from orb_slam2_map_b200 import capi, synth

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "frame_front.npz"))
TUM1 = dict(fx=517.306408, fy=516.469215, cx=318.643040, cy=255.313989,
            k1=0.262383, k2=-0.953104, p1=-0.005358, p2=0.002628, k3=1.163314, bf=40.0)


def _cam(**kw):
    return capi.Camera(**{k: float(v) for k, v in kw.items()})


def test_cvt_gray_and_depth_golden():
    ex = capi.Extractor()
    assert np.array_equal(ex.cvt_gray(G["rgb"], True)[0], G["gray_rgb"])
    assert np.array_equal(ex.cvt_gray(G["rgb"], False)[0], G["gray_bgr"])
    assert np.array_equal(ex.cvt_gray(G["rgba"], True)[0], G["gray_rgba"])
    assert np.array_equal(ex.cvt_gray(G["rgba"], False)[0], G["gray_bgra"])
    rng = np.random.default_rng(4)
    for shape in [(3, 480, 640, 3), (2, 61, 77, 4), (1, 64, 64, 3)]:        # vector path, ragged widths
        img = rng.integers(0, 256, shape, dtype=np.uint8)
        got = ex.cvt_gray(img, True)
        for i in range(shape[0]):
            assert np.array_equal(got[i], O.cvt_gray(img[i], True))
    d16 = rng.integers(0, 65536, (97, 131), dtype=np.uint16)
    f = float(np.float32(1.0) / np.float32(5000.0))
    assert np.array_equal(ex.depth_to_float(d16, f), O.depth_to_float(d16, f))
    ex.close()


def test_undistort_points_and_bounds_golden():
    ex = capi.Extractor()
    for i in range(2):
        c, d = G["cams"][i], G["dists"][i]
        cam = _cam(fx=c[0], fy=c[1], cx=c[2], cy=c[3], k1=d[0], k2=d[1], p1=d[2], p2=d[3], k3=d[4], bf=40.0)
        got = ex.undistort_points(G["pts"], cam)
        assert np.array_equal(got, G[f"undist{i}"]), np.abs(got - G[f"undist{i}"]).max()
        assert np.array_equal(ex.image_bounds(640, 480, cam), O.image_bounds(640, 480, c, d))
    nod = _cam(**{**TUM1, "k1": 0.0})
    assert np.array_equal(ex.undistort_points(G["pts"], nod), G["pts"])
    assert np.array_equal(ex.image_bounds(640, 480, nod), np.array([0, 640, 0, 480], np.float32))
    ex.close()


@pytest.mark.parametrize("channels,rgb,depth16", [(3, True, True), (4, False, False), (1, True, True)])
def test_frames_rgbd_matches_oracle(channels, rgb, depth16):
    w, h, n = 640, 480, 5
    rng = np.random.default_rng(12)
    gray0 = synth.frame_sequence(n, w, h, seed=6)
    if channels == 1:
        color = gray0
    else:
        # a colour image whose luma has the texture of the synthetic frames
        color = np.stack([np.clip(gray0.astype(np.int32) + rng.integers(-20, 21, gray0.shape), 0, 255) for _ in range(channels)], axis=-1).astype(np.uint8)
    if depth16:
        depth = rng.integers(0, 30000, (n, h, w), dtype=np.uint16)
        depth[rng.random(depth.shape) < 0.15] = 0
        factor = float(np.float32(1.0) / np.float32(5000.0))
    else:
        depth = (rng.random((n, h, w)) * 6).astype(np.float32)
        depth[rng.random(depth.shape) < 0.15] = 0
        factor = 1.0
    cam = _cam(**TUM1)
    camv = np.array([TUM1[k] for k in ("fx", "fy", "cx", "cy")], np.float32)
    dist = np.array([TUM1[k] for k in ("k1", "k2", "p1", "p2", "k3")], np.float32)
    ex = capi.Extractor(max_width=w, max_height=h, chunk_frames=2, result_slots=n)
    ex.frames_rgbd(color, depth, cam, rgb_order=rgb, depth_factor=factor)
    kps, desc, counts = ex.download(0, n)
    kun, ur, z, bounds = ex.download_frame_front(0, n)
    assert np.array_equal(bounds, O.image_bounds(w, h, camv, dist))
    orc = O.Extractor()
    frames = []
    for i in range(n):
        gray = color[i] if channels == 1 else O.cvt_gray(color[i], rgb)
        ko, do = orc.extract(gray)
        c = counts[i]
        assert c == len(ko) and kps[i, :c].tobytes() == ko.tobytes() and np.array_equal(desc[i, :c], do)
        xy = np.stack([ko["x"], ko["y"]], axis=1)
        xyu = O.undistort_points(xy, camv, dist)
        dflt = O.depth_to_float(depth[i], factor) if depth16 else depth[i]
        ur_o, z_o = O.stereo_from_rgbd(xy, xyu, dflt, TUM1["bf"])
        assert np.array_equal(kun[i, :c]["x"], xyu[:, 0]) and np.array_equal(kun[i, :c]["y"], xyu[:, 1])
        for f in ("size", "angle", "response", "octave", "class_id"):
            assert np.array_equal(kun[i, :c][f], ko[f])
        assert np.array_equal(ur[i, :c], ur_o) and np.array_equal(z[i, :c], z_o)
        assert (ur[i, c:] == -1).all() and (z[i, c:] == -1).all()
        kou = ko.copy(); kou["x"] = xyu[:, 0]; kou["y"] = xyu[:, 1]
        frames.append(O.FrameData(kou, do, w, h, bounds=tuple(float(b) for b in bounds)))
    # the matcher now runs on mvKeysUn with the undistorted image bounds
    m12, nm, _ = ex.search_init_pairs(0, n - 1)
    for i in range(n - 1):
        f1, f2 = frames[i], frames[i + 1]
        prev = np.stack([f1.kps["x"], f1.kps["y"]], axis=1)
        n_o, m_o, _, _ = O.search_for_initialization(f1, f2, prev, 100, 0.9, True)
        assert nm[i] == n_o and np.array_equal(m12[i, :counts[i]], m_o), f"pair {i}"
    ex.close()


def test_generate_point_clouds_batch():
    from test_oracle_frame import make_cloud_case
    cam = capi.Camera(517.3, 516.5, 318.6, 255.3, 0, 0, 0, 0, 0, 40.0)
    cases = [make_cloud_case(s) for s in range(5)]
    depth = np.stack([c[0] for c in cases]); bgr = np.stack([c[1] for c in cases]); pose = np.stack([c[2] for c in cases])
    depth[4] = 0                                                   # an empty cloud
    for P, Cc in [(pose, bgr), (None, bgr), (pose, None)]:
        got = capi.generate_point_clouds(depth, Cc, cam, P)
        for f in range(len(cases)):
            exp = O.generate_point_cloud(depth[f], None if Cc is None else Cc[f], 517.3, 516.5, 318.6, 255.3,
                                         pose_inv=None if P is None else P[f])
            assert len(got[f]) == len(exp) and got[f].tobytes() == exp.tobytes(), f"frame {f}"
    assert len(got[4]) == 0 and len(got[0]) > 10000
    # odd geometry, step 1 and 2, more sampled rows than threads per pass
    rng = np.random.default_rng(9)
    d = rng.uniform(0, 11, (2, 1100, 37)).astype(np.float32)
    for step in (1, 2, 5):
        got = capi.generate_point_clouds(d, None, cam, pose[:2], step=step)
        for f in range(2):
            exp = O.generate_point_cloud(d[f], None, 517.3, 516.5, 318.6, 255.3, pose_inv=pose[f], step=step)
            assert got[f].tobytes() == exp.tobytes()
    with pytest.raises(capi.OrbxError):
        capi.generate_point_clouds(depth, bgr, cam, pose, capacity=100)


def test_pack_map_records():
    img = synth.frame_sequence(1, 640, 480, seed=12)[0]
    ex = capi.Extractor(max_width=640, max_height=480, chunk_frames=2)
    k, d = ex.extract(img)
    n = len(k)
    assert n > 500
    rng = np.random.default_rng(1)
    mp = rng.integers(0, 1 << 40, n).astype(np.uint64); mp[::4] = np.uint64(0xFFFFFFFFFFFFFFFF)
    got = ex.pack_map_records(0, mp)
    assert got.shape == (n, 64) and got.tobytes() == O.write_keyframe_features(k[:n], d[:n], mp).tobytes()
    assert ex.pack_map_records(0).tobytes() == O.write_keyframe_features(k[:n], d[:n], None).tobytes()


@pytest.mark.parametrize("channels,rgb,depth16", [(3, True, True), (4, False, False), (1, True, True)])
def test_frame_rgbd_single_call_graph_replay(channels, rgb, depth16):
    """orbx_frame_rgbd, the one-frame latency form of the RGB-D front (Tracking::GrabImageRGBD + Frame::Frame): call after
    call on ONE handle — first call of a shape captures the extractor's graph, later calls replay it — every call bit-exact
    against the oracle chain cvtColor -> ORBextractor -> undistortPoints -> ComputeStereoFromRGBD, with strided sources, a
    page-locked source, a shape change and back, and the matcher on the frame it leaves in slot 0."""
    import torch
    w, h, n = 640, 480, 4
    rng = np.random.default_rng(21)
    gray0 = synth.frame_sequence(n, w, h, seed=9)
    color = gray0 if channels == 1 else np.stack([np.clip(gray0.astype(np.int32) + rng.integers(-20, 21, gray0.shape), 0, 255)
                                                   for _ in range(channels)], axis=-1).astype(np.uint8)
    if depth16:
        depth = rng.integers(0, 30000, (n, h, w), dtype=np.uint16); factor = float(np.float32(1.0) / np.float32(5000.0))
    else:
        depth = (rng.random((n, h, w)) * 6).astype(np.float32); factor = 1.0
    depth[rng.random(depth.shape) < 0.15] = 0
    cam = _cam(**TUM1)
    camv = np.array([TUM1[k] for k in ("fx", "fy", "cx", "cy")], np.float32)
    dist = np.array([TUM1[k] for k in ("k1", "k2", "p1", "p2", "k3")], np.float32)
    orc = O.Extractor()
    ex = capi.Extractor(max_width=752, max_height=480, chunk_frames=2, result_slots=2)

    def check_one(col, dep):
        r = ex.frame_rgbd(col, dep, cam, rgb_order=rgb, depth_factor=factor)
        hh, ww = col.shape[:2]
        gray = np.ascontiguousarray(col) if col.ndim == 2 else O.cvt_gray(np.ascontiguousarray(col), rgb)
        ko, do = orc.extract(gray)
        assert r["kps"].tobytes() == ko.tobytes() and np.array_equal(r["desc"], do)
        xy = np.stack([ko["x"], ko["y"]], axis=1)
        xyu = O.undistort_points(xy, camv, dist)
        dflt = O.depth_to_float(np.ascontiguousarray(dep), factor) if depth16 else np.ascontiguousarray(dep)
        ur_o, z_o = O.stereo_from_rgbd(xy, xyu, dflt, TUM1["bf"])
        assert np.array_equal(r["keys_un"]["x"], xyu[:, 0]) and np.array_equal(r["keys_un"]["y"], xyu[:, 1])
        for f in ("size", "angle", "response", "octave", "class_id"):
            assert np.array_equal(r["keys_un"][f], ko[f])
        assert np.array_equal(r["u_right"], ur_o) and np.array_equal(r["depth"], z_o)
        assert np.array_equal(r["bounds"], O.image_bounds(ww, hh, camv, dist))
        return r

    for i in range(n):                                           # call 0 captures, the others replay
        check_one(color[i], depth[i])
    wide = np.zeros((h, w + 60) + color.shape[3:], np.uint8); wide[:, :w] = color[1]; wide[:, w:] = 255      # padded rows
    dwide = np.zeros((h, w + 12), depth.dtype); dwide[:, :w] = depth[1]
    check_one(wide[:, :w], dwide[:, :w])
    pc = torch.empty(color[2].shape, dtype=torch.uint8, pin_memory=True); pc.copy_(torch.from_numpy(color[2]))
    pd = torch.empty(depth[2].shape, dtype=torch.from_numpy(depth[2]).dtype, pin_memory=True); pd.copy_(torch.from_numpy(depth[2]))
    check_one(pc.numpy(), pd.numpy())                            # page-locked sources: DMA in place
    oc = synth.textured_canvas(752, 480, seed=4)                 # another shape, then back
    ocol = oc if channels == 1 else np.stack([oc] * channels, axis=-1)
    odep = (rng.integers(0, 30000, (480, 752), dtype=np.uint16) if depth16 else (rng.random((480, 752)) * 6).astype(np.float32))
    check_one(ocol, odep)
    r = check_one(color[0], depth[0])
    # the frame is in slot 0 with its grid: a stateless matcher call on the same data must agree with the store's view
    _, _, cnt = ex.download(0, 1, want_kps=False, want_desc=False)
    assert cnt[0] == len(r["kps"])
    ex.close()
