"""RANSAC PnP (OpenCvRansacPnp::solvePnp, src/lib/modular_slam/cv_ransac_pnp.cpp:15-85).

CPU part: oracle/pnp_oracle.py against cv2.solvePnPRansac golden outputs (tests/golden/pnp.npz, made by
tests/golden/make_golden_pnp.py with cv2 4.13.0) and against the generating poses.
GPU part (-m gpu): the CUDA path through the C ABI against the oracle (same draws and formulas: identical winning
hypothesis and inlier mask, pose within 1e-9) and against the cv2 goldens.

Tolerances (parity with OpenCV is by tolerance: different minimal solver and random stream):
  inlier set  identical to cv2's (the scenes separate inliers, noise sigma <= 0.7 px, from outliers at a 5 px threshold)
  rvec, tvec  |delta| <= 1e-6 vs cv2 (both are the least-squares pose over the same inliers; observed ~1e-12)
"""
import os

import numpy as np
import pytest

from conftest import GOLDEN

TOL_CV = 1e-6
TOL_ORACLE = 1e-9


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(GOLDEN, "pnp.npz"))


@pytest.fixture(scope="module")
def PO():
    from oracle import pnp_oracle
    return pnp_oracle


def _cases(golden):
    return range(len(golden["cases"]))


def test_oracle_matches_cv2_goldens(PO, golden, pkg):
    fx, fy, cx, cy = pkg.synth.TUM_K
    for i in _cases(golden):
        ok, r, t, mask, h, _, _ = PO.solve_pnp_ransac(golden[f"obj{i}"], golden[f"img{i}"], fx, fy, cx, cy, seed=7, frame=i)
        assert ok and h >= 0
        assert (mask == golden[f"inl{i}"]).all(), f"case {i}: inlier set differs from cv2's"
        assert np.abs(r - golden[f"rvec{i}"]).max() <= TOL_CV and np.abs(t - golden[f"tvec{i}"]).max() <= TOL_CV, i
        # and the pose is the generating one up to the pixel noise
        assert np.abs(r - golden[f"true_rvec{i}"]).max() < 5e-3 and np.abs(t - golden[f"true_tvec{i}"]).max() < 2e-2


def test_oracle_p3p_recovers_exact_pose(PO):
    rng = np.random.default_rng(3)
    for _ in range(50):
        rv, tv = rng.normal(0, 0.4, 3), rng.normal(0, 0.5, 3)
        R = PO.rodrigues_exp(rv)
        Xc = np.stack([rng.uniform(-1, 1, 3), rng.uniform(-1, 1, 3), rng.uniform(1.5, 5, 3)], 1)
        X = (Xc - tv) @ R
        f = Xc / np.sqrt((Xc * Xc).sum(1))[:, None]
        sols = PO.p3p(X, f)
        assert any(np.abs(Rs - R).max() < 1e-6 and np.abs(ts - tv).max() < 1e-6 for Rs, ts in sols)


def test_oracle_quartic_roots(PO):
    rng = np.random.default_rng(5)
    for _ in range(200):
        roots = np.sort(rng.uniform(-3, 3, 4))
        co = np.poly(roots)  # x^4 ...
        got = np.sort(PO.quartic_real_roots(*co))
        assert len(got) == 4 and np.abs(got - roots).max() < 1e-5
    assert PO.quartic_real_roots(1.0, 0.0, 2.0, 0.0, 1.0) == [] or all(  # (x^2 + 1)^2: no real root
        abs(x * x + 1.0) > 0.5 for x in PO.quartic_real_roots(1.0, 0.0, 2.0, 0.0, 1.0))


def test_oracle_rodrigues_roundtrip(PO):
    rng = np.random.default_rng(9)
    for _ in range(100):
        w = rng.normal(0, 1.0, 3)
        if np.sqrt(w @ w) > 3.0:  # the logarithm returns the rotation's principal vector (angle <= pi)
            w *= 3.0 / np.sqrt(w @ w)
        assert np.abs(PO.rodrigues_log(PO.rodrigues_exp(w)) - w).max() < 1e-9
    assert np.abs(PO.rodrigues_log(np.eye(3))).max() == 0.0


def test_oracle_rejects_degenerate_input(PO, pkg):
    fx, fy, cx, cy = pkg.synth.TUM_K
    X, x, *_ = pkg.synth.pnp_scene(1, 3, 0.0, 0.0)
    assert not PO.solve_pnp_ransac(X, x, fx, fy, cx, cy)[0]  # n < 4
    rng = np.random.default_rng(2)
    X = rng.normal(0, 1, (50, 3)).astype(np.float32)
    x = rng.uniform(0, 480, (50, 2)).astype(np.float32)  # no consistent pose: < 5 inliers
    assert not PO.solve_pnp_ransac(X, x, fx, fy, cx, cy, reproj=0.5)[0]


def test_draw_stream_is_uniform_and_deterministic(PO):
    d = [PO.draw(7, 3, h, j, 1000) for h in range(100) for j in range(4)]
    assert d == [PO.draw(7, 3, h, j, 1000) for h in range(100) for j in range(4)]
    assert min(d) >= 0 and max(d) < 1000 and len(set(d)) > 300
    assert PO.sample4(1, 0, 0, 4) is not None and sorted(PO.sample4(1, 0, 0, 4)) == [0, 1, 2, 3]


# ---------------------------------------------------------------------------------------------------------------------
def _batch(golden, idx, cap):
    F = len(idx)
    obj = np.zeros((F, cap, 3), np.float32)
    img = np.zeros((F, cap, 2), np.float32)
    cnt = np.zeros(F, np.int32)
    for k, i in enumerate(idx):
        n = len(golden[f"obj{i}"])
        obj[k, :n], img[k, :n], cnt[k] = golden[f"obj{i}"], golden[f"img{i}"], n
    return obj, img, cnt


@pytest.mark.gpu
def test_gpu_matches_oracle_and_cv2(pkg, PO, golden):
    fx, fy, cx, cy = pkg.synth.TUM_K
    idx = list(_cases(golden))
    obj, img, cnt = _batch(golden, idx, 1024)
    with pkg.Context(max_width=64, max_height=64, max_batch=1, nfeatures=100) as ctx:
        rt, pose, inl, info = ctx.pnp_ransac_batch(obj, img, cnt, pkg.pnp_params(fx, fy, cx, cy, seed=7))
        assert ctx.launch_count >= 1
    for k, i in enumerate(idx):
        n = cnt[k]
        ok, r, t, mask, h, R, tt = PO.solve_pnp_ransac(golden[f"obj{i}"], golden[f"img{i}"], fx, fy, cx, cy, seed=7, frame=k)
        assert info[k, 0] == 1 and ok
        assert info[k, 2] == h, f"case {i}: winning hypothesis {info[k, 2]} vs oracle {h}"
        assert (inl[k, :n] == mask).all() and not inl[k, n:].any() and info[k, 1] == mask.sum()
        assert np.abs(rt[k, :3] - r).max() <= TOL_ORACLE and np.abs(rt[k, 3:] - t).max() <= TOL_ORACLE
        # cv2.solvePnPRansac
        assert (inl[k, :n] == golden[f"inl{i}"]).all()
        assert np.abs(rt[k, :3] - golden[f"rvec{i}"]).max() <= TOL_CV and np.abs(rt[k, 3:] - golden[f"tvec{i}"]).max() <= TOL_CV
        # the reference's result.pose (cv_ransac_pnp.cpp:65-77): orientation = R^-1, position = -(R^-1 t)
        q, p = pose[k, :4], pose[k, 4:]
        Rq = np.array([[1 - 2 * (q[2] ** 2 + q[3] ** 2), 2 * (q[1] * q[2] - q[0] * q[3]), 2 * (q[1] * q[3] + q[0] * q[2])],
                       [2 * (q[1] * q[2] + q[0] * q[3]), 1 - 2 * (q[1] ** 2 + q[3] ** 2), 2 * (q[2] * q[3] - q[0] * q[1])],
                       [2 * (q[1] * q[3] - q[0] * q[2]), 2 * (q[2] * q[3] + q[0] * q[1]), 1 - 2 * (q[1] ** 2 + q[2] ** 2)]])
        assert np.abs(Rq - R.T).max() < 1e-9 and np.abs(p + R.T @ tt).max() < 1e-9 and abs(q @ q - 1.0) < 1e-12


@pytest.mark.gpu
def test_gpu_hypothesis_scores_match_oracle_at_other_parameters(pkg, PO):
    """non-default iterations / threshold / seed / frame_base, ragged batch, tight capacity"""
    fx, fy, cx, cy = 700.0, 710.0, 300.0, 250.0
    scenes = [pkg.synth.pnp_scene(100 + s, n, of, 0.4, K=(fx, fy, cx, cy)) for s, (n, of) in
              enumerate([(57, 0.3), (500, 0.6), (8, 0.0), (5, 0.0), (200, 0.45)])]
    cap = 500
    obj = np.zeros((len(scenes), cap, 3), np.float32)
    img = np.zeros((len(scenes), cap, 2), np.float32)
    cnt = np.array([len(s[0]) for s in scenes], np.int32)
    for k, s in enumerate(scenes):
        obj[k, :cnt[k]], img[k, :cnt[k]] = s[0], s[1]
    params = pkg.pnp_params(fx, fy, cx, cy, seed=12345, iterations=37, reproj_error=3.0, frame_base=11)
    with pkg.Context(max_width=64, max_height=64, max_batch=1, nfeatures=100) as ctx:
        rt, pose, inl, info = ctx.pnp_ransac_batch(obj, img, cnt, params)
    for k, s in enumerate(scenes):
        ok, r, t, mask, h, _, _ = PO.solve_pnp_ransac(s[0], s[1], fx, fy, cx, cy, iters=37, reproj=3.0, seed=12345, frame=11 + k)
        assert bool(info[k, 0]) == ok, k
        if ok:
            assert info[k, 2] == h and (inl[k, :cnt[k]] == mask).all()
            assert np.abs(rt[k, :3] - r).max() <= 1e-8 and np.abs(rt[k, 3:] - t).max() <= 1e-8
            assert np.abs(rt[k, :3] - s[2]).max() < 2e-2 and np.abs(rt[k, 3:] - s[3]).max() < 5e-2


@pytest.mark.gpu
def test_gpu_degenerate_inputs(pkg, PO):
    fx, fy, cx, cy = pkg.synth.TUM_K
    rng = np.random.default_rng(2)
    cap = 64
    obj = np.zeros((4, cap, 3), np.float32)
    img = np.zeros((4, cap, 2), np.float32)
    cnt = np.array([0, 3, 50, 4], np.int32)
    X3, x3, *_ = pkg.synth.pnp_scene(1, 3, 0.0, 0.0)
    obj[1, :3], img[1, :3] = X3, x3
    obj[2, :50] = rng.normal(0, 1, (50, 3))
    img[2, :50] = rng.uniform(0, 480, (50, 2))
    X4, x4, rv4, tv4, _ = pkg.synth.pnp_scene(4, 4, 0.0, 0.0)
    obj[3, :4], img[3, :4] = X4, x4
    params = pkg.pnp_params(fx, fy, cx, cy, reproj_error=0.5)
    with pkg.Context(max_width=64, max_height=64, max_batch=1, nfeatures=100) as ctx:
        rt, pose, inl, info = ctx.pnp_ransac_batch(obj, img, cnt, params)
        # invalid arguments are refused, not computed
        bad = pkg.pnp_params(fx, fy, cx, cy, iterations=0)
        with pytest.raises(pkg.MslamB200Error):
            ctx.pnp_ransac_batch(obj, img, cnt, bad)
        bad = pkg.pnp_params(0.0, fy, cx, cy)
        with pytest.raises(pkg.MslamB200Error):
            ctx.pnp_ransac_batch(obj, img, cnt, bad)
    assert (info[:3, 0] == 0).all() and not inl[:3].any() and (rt[:3] == 0).all()
    assert (pose[:3, 0] == 1).all() and (pose[:3, 1:] == 0).all()
    ok, r, t, mask, h, _, _ = PO.solve_pnp_ransac(X4, x4, fx, fy, cx, cy, reproj=0.5, frame=3)
    assert bool(info[3, 0]) == ok
    if ok:  # four exact points: P3P + the 4th as the check -> all four inliers, the generating pose
        assert info[3, 1] == 4 and np.abs(rt[3, :3] - rv4).max() < 1e-3 and np.abs(rt[3, 3:] - tv4).max() < 1e-3


@pytest.mark.gpu
def test_gpu_batch_is_order_independent_and_repeatable(pkg, golden):
    """frame k of a batch with frame_base b == frame 0 of a single-frame call with frame_base b + k; two runs agree bit for bit"""
    fx, fy, cx, cy = pkg.synth.TUM_K
    obj, img, cnt = _batch(golden, [0, 4, 7], 1000)
    with pkg.Context(max_width=64, max_height=64, max_batch=1, nfeatures=100) as ctx:
        a = ctx.pnp_ransac_batch(obj, img, cnt, pkg.pnp_params(fx, fy, cx, cy, seed=3))
        b = ctx.pnp_ransac_batch(obj, img, cnt, pkg.pnp_params(fx, fy, cx, cy, seed=3))
        for x, y in zip(a, b):
            assert x.tobytes() == y.tobytes()
        for k in range(3):
            s = ctx.pnp_ransac_batch(obj[k:k + 1], img[k:k + 1], cnt[k:k + 1], pkg.pnp_params(fx, fy, cx, cy, seed=3, frame_base=k))
            for x, y in zip(a, s):
                assert x[k].tobytes() == y[0].tobytes()


@pytest.mark.gpu
def test_gpu_damping_retries_match_oracle(pkg, PO):
    """noisy small scenes where the first damping trial of an iteration is rejected: the kernel evaluates trials 2..8
    together (warp per damping) and must take the one the oracle's sequential loop takes"""
    fx, fy, cx, cy = pkg.synth.TUM_K
    spec = [(0, 30, 0.2, 3.0), (0, 12, 0.0, 4.0), (0, 60, 0.4, 2.5), (1, 30, 0.2, 3.0), (1, 12, 0.0, 4.0), (1, 60, 0.4, 2.5)]
    scenes = [pkg.synth.pnp_scene(s, n, o, noise) for s, n, o, noise in spec]
    cap = 64
    obj, img = np.zeros((len(spec), cap, 3), np.float32), np.zeros((len(spec), cap, 2), np.float32)
    cnt = np.array([n for _, n, _, _ in spec], np.int32)
    for k, sc in enumerate(scenes):
        obj[k, :cnt[k]], img[k, :cnt[k]] = sc[0], sc[1]
    with pkg.Context(max_width=64, max_height=64, max_batch=1, nfeatures=100) as ctx:
        rt, pose, inl, info = ctx.pnp_ransac_batch(obj, img, cnt, pkg.pnp_params(fx, fy, cx, cy, seed=7))
    retried = 0
    for k, sc in enumerate(scenes):
        trace = []
        ok, r, t, mask, h, _, _ = PO.solve_pnp_ransac(sc[0], sc[1], fx, fy, cx, cy, seed=7, frame=k, trace=trace)
        assert bool(info[k, 0]) == ok
        if not ok:
            continue
        retried += any(tr > 0 for _, tr, _ in trace)
        assert info[k, 2] == h and (inl[k, :cnt[k]] == mask).all()
        # (the iteration count is not compared: these late trials are accepted or rejected on the last bits of the cost,
        # where the kernel's reduction tree and numpy's summation order may differ; the pose is what must agree)
        assert np.abs(rt[k, :3] - r).max() <= TOL_ORACLE and np.abs(rt[k, 3:] - t).max() <= TOL_ORACLE, k
    assert retried >= 3, "the scenes no longer exercise the damping retries"


@pytest.mark.gpu
def test_gpu_full_size_batch_is_position_independent(pkg):
    """BASELINE-sized batch (64 frames x 1000 correspondences, 50 % outliers): every frame is solved, close to the generating
    pose, and a frame's result does not depend on where it sits in a batch (frame k alone with frame_base = k gives the
    same bytes) -- the size-independent property of a launch that maps one two-CTA cluster to each frame"""
    fx, fy, cx, cy = pkg.synth.TUM_K
    F, n = 64, 1000
    scenes = [pkg.synth.pnp_scene(5000 + i, n, 0.5, 0.7) for i in range(F)]
    obj, img = np.stack([s[0] for s in scenes]), np.stack([s[1] for s in scenes])
    cnt = np.full(F, n, np.int32)
    with pkg.Context(max_width=64, max_height=64, max_batch=1, nfeatures=100) as ctx:
        rt, pose, inl, info = ctx.pnp_ransac_batch(obj, img, cnt, pkg.pnp_params(fx, fy, cx, cy, seed=3))
        assert (info[:, 0] == 1).all() and (info[:, 1] > 0.4 * n).all() and (info[:, 1] == inl.sum(axis=1)).all()
        for k, s in enumerate(scenes):
            assert np.abs(rt[k, :3] - s[2]).max() < 5e-3 and np.abs(rt[k, 3:] - s[3]).max() < 2e-2, k
        for k in (0, 5, 63):
            rt1, pose1, inl1, info1 = ctx.pnp_ransac_batch(obj[k:k + 1], img[k:k + 1], cnt[k:k + 1],
                                                           pkg.pnp_params(fx, fy, cx, cy, seed=3, frame_base=k))
            assert rt1.tobytes() == rt[k:k + 1].tobytes() and pose1.tobytes() == pose[k:k + 1].tobytes()
            assert (inl1 == inl[k:k + 1]).all() and (info1 == info[k:k + 1]).all()
