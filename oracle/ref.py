"""ctypes wrapper round oracle/_ref/libmslam_ref.so -- the REFERENCE's own hot-path sources (frame.cpp, orb_feature.cpp,
distributed_cv_feature.cpp, the back-projection functions of rgbd_feature_frontend.cpp) compiled unchanged by
`make -C oracle _ref` against the stand-in headers of oracle/ref_shim/.  TEST INFRASTRUCTURE ONLY (tests/, and bench.py's
cpu_baseline / --impl reference legs); the product never loads it.

The .so is built in the build container (where /root/reference exists), is git-ignored, and travels to the GPU box with
the snapshot; available() says whether it is there.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from . import oracle as O

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_ref", "libmslam_ref.so")
REFERENCE = os.environ.get("MSLAM_REFERENCE", "/root/reference")
_lib = None


def build():
    """(re)build when the reference tree is present; otherwise use what travelled"""
    if os.path.isdir(os.path.join(REFERENCE, "src", "lib", "modular_slam")):
        subprocess.check_call(["make", "-C", _HERE, "-s", "_ref", f"REF={REFERENCE}"])
    return _SO


def available():
    try:
        lib()
        return True
    except (OSError, subprocess.CalledProcessError):
        return False


def lib():
    global _lib
    if _lib is None:
        O.lib()  # liborb_oracle.so first: the shimmed cv:: primitives forward to it
        if not os.path.exists(_SO):
            build()
        _lib = C.CDLL(_SO)
        for f in ("ref_dist_cos", "ref_dist_sin"):
            getattr(_lib, f).restype = C.c_float
            getattr(_lib, f).argtypes = [C.c_float]
        _lib.ref_dist_ic_angle.restype = C.c_float
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def layout():
    """sizeof(Keypoint), sizeof(OrbKeypoint), offsetof(OrbKeypoint, descriptor), sizeof(DescriptorMatch)"""
    out = np.zeros(4, np.int32)
    lib().ref_layout(_p(out))
    return out.tolist()


def gray(rgb):
    """mslam::toGrayScale (frame.cpp:6-27)"""
    rgb = np.ascontiguousarray(rgb, np.uint8)
    H, W, _ = rgb.shape
    out = np.zeros((H, W), np.uint8)
    lib().ref_gray(_p(rgb), W, H, _p(out))
    return out


def dist_detect(rgb, cap=32768):
    """DistributedOrbOpenCvDetector::detect -> (ids u64, xy [n, 2] f64, descriptors [n, 32])"""
    rgb = np.ascontiguousarray(rgb, np.uint8)
    H, W, _ = rgb.shape
    xy = np.zeros((cap, 2), np.float64)
    ids = np.zeros(cap, np.uint64)
    desc = np.zeros((cap, 32), np.uint8)
    n = lib().ref_dist_detect(_p(rgb), W, H, _p(xy), _p(ids), _p(desc), cap)
    assert n <= cap
    return ids[:n], xy[:n], desc[:n]


def dist_extract(gray_img, cap=32768):
    """OrbExtractorPimpl::extract -> (keypoints[KP_DTYPE] with every cv::KeyPoint field, descriptors)"""
    g = np.ascontiguousarray(gray_img, np.uint8)
    H, W = g.shape
    kps = np.zeros(cap, O.KP_DTYPE)
    desc = np.zeros((cap, 32), np.uint8)
    n = lib().ref_dist_extract(_p(g), W, H, _p(kps), _p(desc), cap)
    assert n <= cap
    return kps[:n], desc[:n]


def dist_tree(xys, min_x, max_x, min_y, max_y, scale_factor):
    """distribute_keypoints_via_tree on (x, y, response) int triples -> the survivors in list order"""
    xys = np.ascontiguousarray(xys, np.int32).reshape(-1, 3)
    out = np.zeros((max(len(xys), 1), 3), np.int32)
    n = lib().ref_dist_tree(_p(xys), len(xys), min_x, max_x, min_y, max_y, C.c_float(scale_factor), _p(out), len(out))
    return out[:n]


def dist_cos(v):
    return lib().ref_dist_cos(C.c_float(v))


def dist_sin(v):
    return lib().ref_dist_sin(C.c_float(v))


def dist_umax():
    out = np.zeros(16, np.int32)
    lib().ref_dist_umax(_p(out))
    return out


def dist_scale_factors():
    out = np.zeros(8, np.float32)
    lib().ref_dist_scale_factors(_p(out))
    return out


def dist_descriptor(blurred, cx, cy, angle_deg):
    b = np.ascontiguousarray(blurred, np.uint8)
    out = np.zeros(32, np.uint8)
    lib().ref_dist_descriptor(_p(b), b.shape[1], b.shape[0], cx, cy, C.c_float(angle_deg), _p(out))
    return out


def dist_ic_angle(img, cx, cy):
    g = np.ascontiguousarray(img, np.uint8)
    return lib().ref_dist_ic_angle(_p(g), g.shape[1], g.shape[0], cx, cy)


def orb_detect(rgb, cap=16384):
    """OrbOpenCvDetector::detect -> (xy [n, 2] f64, descriptors); cv::ORB itself is the cv2-pinned oracle (see the shim)"""
    rgb = np.ascontiguousarray(rgb, np.uint8)
    H, W, _ = rgb.shape
    xy = np.zeros((cap, 2), np.float64)
    desc = np.zeros((cap, 32), np.uint8)
    n = lib().ref_orb_detect(_p(rgb), W, H, _p(xy), _p(desc), cap)
    assert n <= cap
    return xy[:n], desc[:n]


def orb_match(from_desc, to_desc):
    """OrbOpenCvMatcher::match(from, to) on OrbKeypoint records -> (fromIndex, toIndex) u64"""
    def recs(d):
        r = np.zeros((len(d), 64), np.uint8)
        r[:, 32:] = d
        return r
    f, t = recs(from_desc), recs(to_desc)
    cap = max(len(t), 1)
    fi = np.zeros(cap, np.uint64)
    ti = np.zeros(cap, np.uint64)
    n = lib().ref_orb_match(_p(f), len(f), _p(t), len(t), _p(fi), _p(ti), cap)
    return fi[:n], ti[:n]


def backproject(xy, depth, cx, cy, fx, fy, factor):
    """pointsFromRgbdKeypoints (rgbd_feature_frontend.cpp:119-138) -> (xyz [n, 3] f64, valid [n] u8)"""
    xy = np.ascontiguousarray(xy, np.float64).reshape(-1, 2)
    depth = np.ascontiguousarray(depth, np.uint16)
    n = len(xy)
    xyz = np.zeros((max(n, 1), 3), np.float64)
    valid = np.zeros(max(n, 1), np.uint8)
    lib().ref_backproject(_p(xy), n, _p(depth), depth.shape[1], depth.shape[0], C.c_double(cx), C.c_double(cy), C.c_double(fx),
                          C.c_double(fy), C.c_float(factor), _p(xyz), _p(valid))
    return xyz[:n], valid[:n]
