"""depth back-projection of matched keypoints (rgbd_feature_frontend.cpp:101-138).
CPU part: the oracle against an independent numpy float64 evaluation on the reference's bundled depth frame.
GPU part (-m gpu): the CUDA path through the C ABI, bit-exact against the oracle."""
import os

import numpy as np
import pytest

from conftest import GOLDEN


@pytest.fixture(scope="module")
def depth():
    return np.load(os.path.join(GOLDEN, "depth.npz"))


def _xy_from_golden(bundled, i=0):
    kp = bundled[f"kp{i}"]
    return np.stack([kp["x"].astype(np.float64), kp["y"].astype(np.float64)], axis=1)  # (double)pt, orb_feature.cpp:57


def _numpy_backproject(xy, depth, cam):
    xi, yi = xy[:, 0].astype(np.int64), xy[:, 1].astype(np.int64)  # truncation
    d = depth[yi, xi].astype(np.int32).astype(np.float32) * np.float32(cam["factor"])
    ok = d > np.finfo(np.float32).eps
    z = d.astype(np.float64)
    x = (xy[:, 0] - cam["cx"]) * z * (1.0 / cam["fx"])
    y = (xy[:, 1] - cam["cy"]) * z * (1.0 / cam["fy"])
    out = np.stack([x, y, z], axis=1)
    out[~ok] = 0.0
    return out, ok.astype(np.uint8)


def test_oracle_matches_independent_evaluation(oracle, bundled, depth):
    cam = oracle.TUM_CAMERA
    xy = _xy_from_golden(bundled)
    xyz, valid = oracle.backproject(xy, depth["depth0"], **cam)
    ref, ok = _numpy_backproject(xy, depth["depth0"], cam)
    assert (valid == ok).all() and xyz.tobytes() == ref.tobytes()
    assert 0 < valid.sum() < len(valid) or valid.all()  # the bundled frame has holes (5 % zero depth)
    assert np.isclose(xyz[valid == 1][:, 2], depth["depth0"][xy[valid == 1][:, 1].astype(int), xy[valid == 1][:, 0].astype(int)] / 5000.0).all()


def test_oracle_strided_records_and_edge_cases(oracle, depth):
    cam = oracle.TUM_CAMERA
    rec = np.zeros(5, np.dtype({"names": ["id", "coordinates", "descriptor"], "formats": ["<u8", ("<f8", (2,)), ("u1", (32,))],
                                "offsets": [0, 16, 32], "itemsize": 64}))  # OrbKeypoint layout
    d = np.zeros((4, 6), np.uint16)
    d[1, 2], d[3, 5] = 5000, 1
    rec["coordinates"] = [[2.9, 1.9], [5.0, 3.0], [0.0, 0.0], [6.0, 1.0], [-0.5, 1.0]]
    xyz, valid = oracle.backproject(rec["coordinates"], d, **cam)
    assert valid.tolist() == [1, 1, 0, 0, 1] or valid.tolist() == [1, 1, 0, 0, 0]
    assert xyz[0, 2] == 1.0 and valid[2] == 0 and valid[3] == 0  # zero depth -> nullopt; x = 6 is outside the frame
    assert xyz[0, 0] == (2.9 - 319.5) * 1.0 * (1.0 / 525.0)
    # (int)(-0.5) == 0: truncation toward zero keeps the point inside, exactly as Eigen's cast<int>() does
    assert valid[4] == 0 and d[1, 0] == 0
    e, v = oracle.backproject(np.zeros((0, 2)), d, **cam)
    assert e.shape == (0, 3) and v.shape == (0,)


@pytest.mark.gpu
def test_gpu_backproject_bit_exact(pkg, oracle, bundled, depth):
    cam = oracle.TUM_CAMERA
    c = pkg.Camera(cx=cam["cx"], cy=cam["cy"], fx=cam["fx"], fy=cam["fy"], factor=float(cam["factor"]))
    with pkg.Context() as ctx:
        for i in (0, 1):
            xy = _xy_from_golden(bundled, i)
            xyz, valid = ctx.backproject(xy, depth[f"depth{i}"], c)
            rx, rv = oracle.backproject(xy, depth[f"depth{i}"], **cam)
            assert (valid == rv).all() and xyz.tobytes() == rx.tobytes()
        # arbitrary (not float-representable) coordinates, strided 64-byte records, out-of-frame points
        rng = np.random.default_rng(4)
        rec = np.zeros(3000, np.dtype({"names": ["id", "coordinates"], "formats": ["<u8", ("<f8", (2,))], "offsets": [0, 16],
                                       "itemsize": 64}))
        rec["coordinates"] = rng.uniform(-3.0, 645.0, (3000, 2))
        xyz, valid = ctx.backproject(rec["coordinates"], depth["depth0"], c)
        rx, rv = oracle.backproject(rec["coordinates"], depth["depth0"], **cam)
        assert (valid == rv).all() and xyz.tobytes() == rx.tobytes() and 0 < rv.sum() < 3000
        e, v = ctx.backproject(np.zeros((0, 2)), depth["depth0"], c)
        assert e.shape == (0, 3)


@pytest.mark.gpu
def test_gpu_match_to_3d_association_on_device(pkg, oracle, bundled, depth):
    """detect -> match -> back-project the matched keypoints without leaving the device (track(), :292-293)"""
    import torch
    cam = oracle.TUM_CAMERA
    c = pkg.Camera(cx=cam["cx"], cy=cam["cy"], fx=cam["fx"], fy=cam["fy"], factor=float(cam["factor"]))
    with pkg.Context(max_batch=2) as ctx:
        cap = ctx.kp_cap
        frames = np.stack([bundled["bgr1"], bundled["bgr0"]])  # pair 0: current = frame 1, pair 1: current = frame 0
        d_fr = torch.from_numpy(frames).cuda()
        d_kp = torch.zeros((2, cap, 6), dtype=torch.float32, device="cuda")
        d_desc = torch.zeros((2, cap, 32), dtype=torch.uint8, device="cuda")
        d_cnt = torch.zeros(2, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        ctx.detect_batch_dev(d_fr, 2, 640, 480, d_kp, d_desc, cap, d_cnt)
        d_prev = torch.flip(d_desc, [0]).contiguous()
        d_pcnt = torch.flip(d_cnt, [0]).contiguous()
        out = [torch.zeros((2, cap), dtype=torch.int32, device="cuda") for _ in range(3)]
        d_n = torch.zeros(2, dtype=torch.int32, device="cuda")
        ctx.synchronize()
        ctx.match_batch_dev(0, d_desc, d_cnt, d_prev, d_pcnt, 2, cap, out[0], out[1], out[2], d_n)  # from = current frame
        d_depth = torch.from_numpy(np.stack([depth["depth1"], depth["depth0"]]).astype(np.int16)).cuda()
        d_xyz = torch.zeros((2, cap, 3), dtype=torch.float64, device="cuda")
        d_val = torch.zeros((2, cap), dtype=torch.uint8, device="cuda")
        ctx.backproject_dev(d_kp, cap, out[0], d_n, 2, cap, d_depth, 640, 480, c, d_xyz, d_val)
        ctx.synchronize()
        kp, n, fi = d_kp.cpu().numpy(), d_n.cpu().numpy(), out[0].cpu().numpy()
        for p, dname in ((0, "depth1"), (1, "depth0")):
            assert n[p] > 300
            xy = kp[p, fi[p, :n[p]], :2].astype(np.float64)
            rx, rv = oracle.backproject(xy, depth[dname], **cam)
            assert (d_val[p, :n[p]].cpu().numpy() == rv).all()
            assert d_xyz[p, :n[p]].cpu().numpy().tobytes() == rx.tobytes()
