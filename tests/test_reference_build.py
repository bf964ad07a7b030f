"""The checker checked by the reference itself: oracle/_ref/libmslam_ref.so is the REFERENCE's own code for the path --
frame.cpp, orb_feature.cpp, distributed_cv_feature.cpp and the back-projection functions of rgbd_feature_frontend.cpp,
compiled unchanged from /root/reference by `make -C oracle _ref` (stand-in Eigen / OpenCV containers; the cv:: arithmetic
primitives forward to the cv2-4.13.0-pinned C oracle, see oracle/ref_shim/mslam_cv_shim.hpp; the reference pins 4.8.1).

CPU half (-m "not gpu"): the C restatements in oracle/ (orb_oracle.c, dist_orb_oracle.c) must equal the reference build
bit for bit -- gray, the whole second detector in the reference's output order with every cv::KeyPoint field, its quad-tree,
polynomial cos/sin, u_max, scale factors, IC angle, descriptor sampling; OrbOpenCvMatcher::match; pointsFromRgbdKeypoints.
GPU half (-m gpu): the CUDA path, through the C ABI, against the reference build directly (no restatement in between).

The .so is built in the build container and travels with the snapshot; nothing here reads /root/reference at run time.
"""
import os

import numpy as np
import pytest

from conftest import GOLDEN


@pytest.fixture(scope="module")
def ref():
    from oracle import ref as R
    if not R.available():
        pytest.skip("oracle/_ref/libmslam_ref.so not built (needs the reference tree at build time)")
    return R


def _dist_frames(pkg, bundled):
    """(name, BGR frame) cases: bundled, the three configured sizes, odd / portrait sizes, noise, low contrast, flat"""
    cases = [("bundled0", bundled["bgr0"]), ("bundled1", bundled["bgr1"])]
    for W, H in ((640, 480), (1280, 720), (1920, 1080), (311, 237), (480, 640), (333, 500), (160, 140)):
        cases.append((f"synth{W}x{H}", pkg.synth.frame_batch(W, H, 1, 0, seed=W + H)[0][0]))
    rng = np.random.default_rng(11)
    cases.append(("noise", rng.integers(0, 256, (480, 640, 3), dtype=np.uint8)))
    g = (120 + np.random.default_rng(5).integers(0, 14, (480, 640))).astype(np.uint8)  # contrast 7..20: threshold fallback
    cases.append(("lowcontrast", np.repeat(g[:, :, None], 3, axis=2)))
    cases.append(("flat", np.full((240, 320, 3), 77, np.uint8)))
    return cases


# ---- a1: toGrayScale ----------------------------------------------------------------------------------------------------
def test_layouts_of_the_reference_types(ref):
    assert ref.layout() == [32, 64, 32, 16]  # Keypoint, OrbKeypoint, offsetof(descriptor), DescriptorMatch


def test_gray_oracle_equals_frame_cpp(ref, oracle, bundled):
    assert (oracle.gray(bundled["bgr0"]) == ref.gray(bundled["bgr0"])).all()
    # every (c0, c1, c2) with two channels swept exhaustively at a few values of the third, plus the saturating corner
    a = np.arange(256, dtype=np.uint8)
    for c in (0, 1, 37, 128, 200, 254, 255):
        img = np.stack(np.broadcast_arrays(a[:, None], a[None, :], np.uint8(c)), -1)
        for perm in ((0, 1, 2), (2, 0, 1), (1, 2, 0)):
            x = np.ascontiguousarray(img[..., perm])
            assert (oracle.gray(x) == ref.gray(x)).all()


# ---- f1: the second detector ----------------------------------------------------------------------------------------------
def test_dist_constants_equal_the_reference(ref, oracle):
    assert (oracle.dist_umax() == ref.dist_umax()).all()
    _, _, s = oracle.dist_level_geometry(640, 480)
    assert s.tobytes() == ref.dist_scale_factors().tobytes()


def test_dist_cos_sin_equal_the_reference(ref, oracle):
    rng = np.random.default_rng(3)
    vals = np.concatenate([np.linspace(-7, 7, 20001), rng.uniform(-400, 400, 20000), np.deg2rad(np.arange(0, 361, 0.25)),
                           [0.0, np.pi / 2, np.pi, 1.5 * np.pi, 2 * np.pi, -np.pi, 1e-30, -1e-30]]).astype(np.float32)
    for v in vals:
        assert np.float32(oracle.dist_cos(float(v))).tobytes() == np.float32(ref.dist_cos(float(v))).tobytes(), v
        assert np.float32(oracle.dist_sin(float(v))).tobytes() == np.float32(ref.dist_sin(float(v))).tobytes(), v


def test_dist_whole_detector_oracle_equals_the_reference(ref, oracle, pkg, bundled):
    for name, bgr in _dist_frames(pkg, bundled):
        okp, odesc = oracle.dist_detect_rgb(bgr)
        # OrbExtractorPimpl::extract: x, y, size, angle, response, octave of every keypoint, reference order
        rkp, rdesc = ref.dist_extract(ref.gray(bgr))
        assert len(okp) == len(rkp), f"{name}: oracle {len(okp)} keypoints, reference {len(rkp)}"
        assert okp.tobytes() == rkp.tobytes(), f"{name}: keypoint fields differ"
        assert (odesc == rdesc).all(), f"{name}: descriptors differ"
        # the public detect(): ids 0..n-1, coordinates widened to double
        ids, xy, desc = ref.dist_detect(bgr)
        assert (ids == np.arange(len(okp))).all()
        assert (xy[:, 0] == okp["x"].astype(np.float64)).all() and (xy[:, 1] == okp["y"].astype(np.float64)).all()
        assert (desc == odesc).all()
        if name != "flat":
            assert len(okp) > 0


def test_dist_tree_oracle_equals_the_reference(ref, oracle):
    """distribute_keypoints_via_tree on random candidate sets: dense, sparse, clustered, duplicate positions, portrait"""
    rng = np.random.default_rng(17)
    for trial in range(40):
        w, h = [(602, 442), (442, 602), (300, 300), (1242, 682), (90, 70)][trial % 5]
        n = int(rng.integers(1, 4000))
        if trial % 3 == 0:  # clustered: many keypoints fall into one leaf, responses tie
            x = np.clip(rng.normal(w / 2, w / 20, n), 0, w - 1).astype(np.int32)
            y = np.clip(rng.normal(h / 2, h / 20, n), 0, h - 1).astype(np.int32)
        else:
            x, y = rng.integers(0, w, n).astype(np.int32), rng.integers(0, h, n).astype(np.int32)
        xys = np.stack([x, y, rng.integers(7, 60, n).astype(np.int32)], 1)
        sf = float(np.float32(1.2) ** (trial % 8))
        kept, _ = oracle.dist_tree(xys, w + 38, h + 38, sf)  # indices of the survivors, list order
        want = ref.dist_tree(xys, 19, w + 19, 19, h + 19, sf)
        assert len(kept) == len(want) and (xys[kept] == want).all(), f"trial {trial}"


def test_dist_angle_and_descriptor_equal_the_reference(ref, oracle, bundled):
    g = oracle.gray(bundled["bgr0"])
    b = oracle.gauss7_u8(g)
    rng = np.random.default_rng(2)
    for _ in range(300):
        x, y = int(rng.integers(19, 640 - 19)), int(rng.integers(19, 480 - 19))
        a_o = np.float32(oracle.ic_angle(g, x, y))
        assert a_o.tobytes() == np.float32(ref.dist_ic_angle(g, x, y)).tobytes()
        ang = float(rng.uniform(0, 360)) if rng.random() < 0.8 else float(a_o)
        assert (oracle.dist_describe(b, x, y, ang) == ref.dist_descriptor(b, x, y, ang)).all()


# ---- a2 / a4: the OpenCV wrappers of orb_feature.cpp -------------------------------------------------------------------------
def test_orb_detector_wrapper(ref, oracle, bundled):
    """OrbOpenCvDetector::detect keeps (x, y) widened to double and the descriptor, in detectAndCompute's order"""
    xy, desc = ref.orb_detect(bundled["bgr0"])
    okp, od = oracle.orb_detect_rgb(bundled["bgr0"])
    assert len(xy) == 1000 and (xy[:, 0] == okp["x"].astype(np.float64)).all() and (xy[:, 1] == okp["y"].astype(np.float64)).all()
    assert (desc == od).all()


def test_orb_matcher_oracle_equals_the_reference(ref, oracle, bundled, synthetic):
    gb = np.load(os.path.join(GOLDEN, "bench_synth.npz"))
    pairs = [(bundled["desc0"], bundled["desc1"]), (synthetic["tie_b"], synthetic["tie_a"]),
             (gb["640x480_desc0"], gb["640x480_desc1"]), (gb["1280x720_desc0"], gb["1280x720_desc1"]),
             (bundled["desc0"][:2], bundled["desc1"][:7]), (bundled["desc0"][:300], bundled["desc1"])]
    for frm, to in pairs:
        fi, ti, _ = oracle.match_ratio(frm, to)
        rfi, rti = ref.orb_match(frm, to)
        assert len(fi) == len(rfi) and (fi == rfi.astype(np.int64)).all() and (ti == rti.astype(np.int64)).all()
    rfi, rti = ref.orb_match(bundled["desc0"], bundled["desc1"])
    assert len(rfi) == 693  # SURVEY 8(c) golden count, now from the reference's own match()


# ---- f2: pointsFromRgbdKeypoints -----------------------------------------------------------------------------------------------
def _xy_cases(depth):
    rng = np.random.default_rng(8)
    H, W = depth.shape
    xy = np.stack([rng.uniform(0, W - 1, 3000), rng.uniform(0, H - 1, 3000)], 1)
    xy[:200] = np.floor(xy[:200])           # integer coordinates
    xy[200:300] = np.float32(xy[200:300])   # float-valued, as the detectors produce them
    xy[300] = (0.0, 0.0)
    xy[301] = (W - 1, H - 1)
    xy[302] = (W - 0.001, H - 1)            # truncates into the last column
    return xy


def test_backproject_oracle_equals_the_reference(ref, oracle):
    d = np.load(os.path.join(GOLDEN, "depth.npz"))
    cam = oracle.TUM_CAMERA
    for name in ("depth0", "depth1"):
        depth = d[name]
        xy = _xy_cases(depth)
        oxyz, ov = oracle.backproject(xy, depth, **cam)
        rxyz, rv = ref.backproject(xy, depth, **cam)
        assert (ov == rv).all() and 0 < int(rv.sum()) < len(rv)  # the bundled depth frames have holes
        assert oxyz.tobytes() == rxyz.tobytes()
    z = np.zeros((4, 4), np.uint16)
    z[1, 1] = 1  # 1 * factor = 2e-4 > FLT_EPSILON: valid; 0 is not
    oxyz, ov = oracle.backproject(np.array([[1.0, 1.0], [2.0, 2.0]]), z, **cam)
    rxyz, rv = ref.backproject(np.array([[1.0, 1.0], [2.0, 2.0]]), z, **cam)
    assert ov.tolist() == rv.tolist() == [1, 0] and oxyz.tobytes() == rxyz.tobytes()


# ---- the CUDA path against the reference build, no restatement in between ----------------------------------------------------------
@pytest.mark.gpu
def test_gpu_second_detector_equals_the_reference(ref, pkg, bundled):
    for name, bgr in _dist_frames(pkg, bundled):
        H, W, _ = bgr.shape
        rkp, rdesc = ref.dist_extract(ref.gray(bgr))
        with pkg.DistContext(max_width=W, max_height=H, max_batch=1) as ctx:
            kp, desc = ctx.detect(bgr)
        assert len(kp) == len(rkp), f"{name}: {len(kp)} keypoints, reference {len(rkp)}"
        assert kp.tobytes() == rkp.tobytes(), f"{name}: keypoint fields differ from the reference"
        assert (desc == rdesc).all(), f"{name}: descriptors differ from the reference"


@pytest.mark.gpu
def test_gpu_second_detector_frontend_ids_equal_the_reference(ref, pkg, bundled):
    det = pkg.DistributedOrbB200Detector(max_width=640, max_height=480)
    out = det.detect(pkg.RgbFrame(bundled["bgr1"], 640, 480))
    ids, xy, desc = ref.dist_detect(bundled["bgr1"])
    assert (out["id"] == ids).all() and (out["coordinates"] == xy).all() and (out["descriptor"] == desc).all()
    det.ctx.close()


@pytest.mark.gpu
def test_gpu_gray_equals_frame_cpp(ref, pkg, bundled):
    with pkg.Context(max_width=640, max_height=480, max_batch=1, nfeatures=1000) as ctx:
        ctx.detect(bundled["bgr0"])
        got = ctx.debug_read(pkg.STAGE_PYRAMID, 0, 0, 640, 480)
    assert (got == ref.gray(bundled["bgr0"])).all()


@pytest.mark.gpu
@pytest.mark.parametrize("flags_name", ["FLAG_KNN_SCALAR", "FLAG_KNN_TENSOR"])
def test_gpu_matcher_equals_the_reference_match(ref, pkg, bundled, synthetic, flags_name):
    gb = np.load(os.path.join(GOLDEN, "bench_synth.npz"))
    pairs = [(bundled["desc0"], bundled["desc1"]), (synthetic["tie_b"], synthetic["tie_a"]),
             (gb["1280x720_desc0"], gb["1280x720_desc1"]), (gb["1920x1080_desc0"], gb["1920x1080_desc1"])]
    with pkg.Context(max_width=64, max_height=64, max_batch=1, nfeatures=4000, max_descriptors=4000,
                     flags=getattr(pkg, flags_name)) as ctx:
        for frm, to in pairs:
            fi, ti, _ = ctx.match_knn_ratio(frm, to)
            rfi, rti = ref.orb_match(frm, to)
            assert len(fi) == len(rfi) and (fi == rfi).all() and (ti == rti).all()


@pytest.mark.gpu
def test_gpu_backproject_equals_the_reference(ref, oracle, pkg):
    d = np.load(os.path.join(GOLDEN, "depth.npz"))
    cam = oracle.TUM_CAMERA
    with pkg.Context(max_width=640, max_height=480, max_batch=1, nfeatures=1000) as ctx:
        for name in ("depth0", "depth1"):
            xy = _xy_cases(d[name])
            xyz, valid = ctx.backproject(xy, d[name], pkg.Camera(cam["cx"], cam["cy"], cam["fx"], cam["fy"], cam["factor"]))
            rxyz, rv = ref.backproject(xy, d[name], **cam)
            assert (valid == rv).all() and xyz.tobytes() == rxyz.tobytes()
