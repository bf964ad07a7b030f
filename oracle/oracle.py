"""ctypes wrapper round oracle/liborb_oracle.so -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this
module.  The product (modular-slam_b200) never does.  See the header of orb_oracle.c for what the
oracle restates (reference: src/lib/modular_slam/orb_feature.cpp:33-65,84-117; frame.cpp:6-27) and how
it is pinned.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liborb_oracle.so")


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in ("orb_oracle.c", "dist_orb_oracle.c")]
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < max(os.path.getmtime(s) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return _SO


class Params(C.Structure):
    _fields_ = [("nfeatures", C.c_int), ("nlevels", C.c_int), ("fast_threshold", C.c_int), ("edge", C.c_int),
                ("scale_factor", C.c_float)]


class Dumps(C.Structure):
    _fields_ = [("pyr", C.c_void_p), ("blurred", C.c_void_p), ("n_fast", C.c_void_p), ("n_cut1", C.c_void_p)]


KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4")])

_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        try:
            _lib = C.CDLL(_SO)
        except OSError:
            build(force=True)
            _lib = C.CDLL(_SO)
        _lib.orc_harris.restype = C.c_float
        _lib.orc_fast_atan2.restype = C.c_float
        _lib.orc_fast_atan2.argtypes = [C.c_float, C.c_float]
        _lib.orc_ic_angle.restype = C.c_float
        _lib.orc_dist_cos.restype = C.c_float
        _lib.orc_dist_cos.argtypes = [C.c_float]
        _lib.orc_dist_sin.restype = C.c_float
        _lib.orc_dist_sin.argtypes = [C.c_float]
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def default_params(nfeatures=1000, nlevels=8, scale_factor=1.2, fast_threshold=20, edge=31):
    return Params(nfeatures, nlevels, fast_threshold, edge, scale_factor)


def level_geometry(W, H, nlevels=8, scale_factor=1.2):
    w = np.zeros(nlevels, np.int32)
    h = np.zeros(nlevels, np.int32)
    s = np.zeros(nlevels, np.float32)
    lib().orc_level_geometry(W, H, nlevels, C.c_float(scale_factor), _p(w), _p(h), _p(s))
    return w, h, s


def features_per_level(nfeatures=1000, nlevels=8, scale_factor=1.2):
    n = np.zeros(nlevels, np.int32)
    lib().orc_features_per_level(nfeatures, nlevels, C.c_float(scale_factor), _p(n))
    return n


def gray(rgb):
    rgb = np.ascontiguousarray(rgb, np.uint8)
    out = np.empty(rgb.shape[:-1], np.uint8)
    lib().orc_gray(_p(rgb), C.c_long(out.size), _p(out))
    return out


def resize_coeffs(src, dst):
    ofs = np.zeros(dst, np.int32)
    a1 = np.zeros(dst, np.int32)
    lib().orc_resize_coeffs(src, dst, _p(ofs), _p(a1))
    return ofs, a1


def resize_exact(img, dw, dh):
    img = np.ascontiguousarray(img, np.uint8)
    out = np.empty((dh, dw), np.uint8)
    lib().orc_resize_exact(_p(img), img.shape[1], img.shape[0], _p(out), dw, dh)
    return out


def fast_score_map(img, thr=20):
    img = np.ascontiguousarray(img, np.uint8)
    out = np.empty_like(img)
    lib().orc_fast_score_map(_p(img), img.shape[1], img.shape[0], thr, _p(out))
    return out


def fast_nms_border(score, edge=31):
    score = np.ascontiguousarray(score, np.uint8)
    cap = score.size // 4 + 16
    out = np.empty((cap, 3), np.int32)
    n = lib().orc_fast_nms_border(_p(score), score.shape[1], score.shape[0], edge, _p(out), cap)
    return out[:n].copy()


def harris(img, x, y):
    img = np.ascontiguousarray(img, np.uint8)
    return np.float32(lib().orc_harris(_p(img), img.shape[1], int(x), int(y)))


def ic_angle(img, x, y):
    img = np.ascontiguousarray(img, np.uint8)
    return np.float32(lib().orc_ic_angle(_p(img), img.shape[1], int(x), int(y)))


def fast_atan2(y, x):
    return np.float32(lib().orc_fast_atan2(float(y), float(x)))


def blur7(img):
    img = np.ascontiguousarray(img, np.uint8)
    out = np.empty_like(img)
    lib().orc_blur7(_p(img), img.shape[1], img.shape[0], _p(out))
    return out


def describe(blurred, cx, cy, angle_deg):
    blurred = np.ascontiguousarray(blurred, np.uint8)
    out = np.empty(32, np.uint8)
    lib().orc_describe(_p(blurred), blurred.shape[1], int(cx), int(cy), C.c_float(angle_deg), _p(out))
    return out


def orb_detect_gray(gray_img, params=None, cap=None, dumps=False):
    """-> (keypoints[KP_DTYPE], descriptors[N,32]) in canonical (octave, y, x) order.
    dumps=True additionally returns a dict with pyramid levels, blurred levels, n_fast, n_cut1."""
    g = np.ascontiguousarray(gray_img, np.uint8)
    H, W = g.shape
    prm = params or default_params()
    cap = cap or max(4 * prm.nfeatures, 64)
    kps = np.zeros(cap, KP_DTYPE)
    desc = np.zeros((cap, 32), np.uint8)
    d = None
    extra = None
    if dumps:
        w, h, _ = level_geometry(W, H, prm.nlevels, prm.scale_factor)
        tot = int((w.astype(np.int64) * h).sum())
        pyr = np.zeros(tot, np.uint8)
        blr = np.zeros(tot, np.uint8)
        nf = np.zeros(prm.nlevels, np.int32)
        nc = np.zeros(prm.nlevels, np.int32)
        d = Dumps(_p(pyr).value, _p(blr).value, _p(nf).value, _p(nc).value)
        extra = (w, h, pyr, blr, nf, nc)
    n = lib().orc_orb_detect_gray(_p(g), W, H, C.byref(prm), _p(kps), _p(desc), cap, C.byref(d) if d else None)
    if n > cap:
        return orb_detect_gray(g, prm, cap=n, dumps=dumps)
    if not dumps:
        return kps[:n].copy(), desc[:n].copy()
    w, h, pyr, blr, nf, nc = extra
    offs = np.concatenate([[0], np.cumsum(w.astype(np.int64) * h)])
    out = {
        "pyr": [pyr[offs[l]:offs[l + 1]].reshape(h[l], w[l]) for l in range(prm.nlevels)],
        "blurred": [blr[offs[l]:offs[l + 1]].reshape(h[l], w[l]) for l in range(prm.nlevels)],
        "n_fast": nf, "n_cut1": nc,
    }
    return kps[:n].copy(), desc[:n].copy(), out


def orb_detect_rgb(rgb, params=None, cap=None):
    """OrbOpenCvDetector::detect restated (gray conversion in memory order, then ORB)."""
    return orb_detect_gray(gray(rgb), params, cap)


def orb_detect_rgb_batch(frames, params=None, cap=None, nthreads=1):
    frames = np.ascontiguousarray(frames, np.uint8)
    B, H, W, _ = frames.shape
    prm = params or default_params()
    cap = cap or 2 * prm.nfeatures
    kps = np.zeros((B, cap), KP_DTYPE)
    desc = np.zeros((B, cap, 32), np.uint8)
    counts = np.zeros(B, np.int32)
    lib().orc_orb_detect_rgb_batch(_p(frames), B, W, H, C.byref(prm), _p(kps), _p(desc), cap, _p(counts), nthreads)
    assert (counts <= cap).all(), "oracle batch capacity exceeded"
    return kps, desc, counts


def _desc_view(d):
    d = np.asarray(d)
    assert d.dtype == np.uint8 and d.ndim == 2 and d.shape[1] >= 32
    if d.strides[1] != 1:
        d = np.ascontiguousarray(d)
    return d, d.strides[0]


def knn2(query, train):
    q, qs = _desc_view(query)
    t, ts = _desc_view(train)
    nq, nt = q.shape[0], t.shape[0]
    out = np.zeros((4, nq), np.int32)
    lib().orc_knn2(_p(q), nq, qs, _p(t), nt, ts, _p(out[0]), _p(out[1]), _p(out[2]), _p(out[3]))
    return out[0], out[1], out[2], out[3]  # idx0, d0, idx1, d1


def match_ratio(frm, to, ratio=0.7):
    """OrbOpenCvMatcher::match(from, to) -> (fromIndex[int64], toIndex[int64], dist[int32])."""
    f, fs = _desc_view(frm) if len(frm) else (np.zeros((0, 32), np.uint8), 32)
    t, ts = _desc_view(to) if len(to) else (np.zeros((0, 32), np.uint8), 32)
    nt = t.shape[0]
    fi = np.zeros(max(nt, 1), np.int64)
    ti = np.zeros(max(nt, 1), np.int64)
    di = np.zeros(max(nt, 1), np.int32)
    n = lib().orc_match_ratio(_p(f), f.shape[0], fs, _p(t), nt, ts, C.c_double(ratio), _p(fi), _p(ti), _p(di))
    return fi[:n].copy(), ti[:n].copy(), di[:n].copy()


def match_crosscheck(query, train):
    q, qs = _desc_view(query) if len(query) else (np.zeros((0, 32), np.uint8), 32)
    t, ts = _desc_view(train) if len(train) else (np.zeros((0, 32), np.uint8), 32)
    nq = q.shape[0]
    qi = np.zeros(max(nq, 1), np.int32)
    ti = np.zeros(max(nq, 1), np.int32)
    di = np.zeros(max(nq, 1), np.int32)
    n = lib().orc_match_crosscheck(_p(q), nq, qs, _p(t), t.shape[0], ts, _p(qi), _p(ti), _p(di))
    return qi[:n].copy(), ti[:n].copy(), di[:n].copy()


def db_scores(queries, q_counts, db, k_counts, nthreads=1):
    """queries [Q, qcap, 32], db [K, kcap, 32] -> scores [Q, K] (int32)."""
    queries = np.ascontiguousarray(queries, np.uint8)
    db = np.ascontiguousarray(db, np.uint8)
    qc = np.ascontiguousarray(q_counts, np.int32)
    kc = np.ascontiguousarray(k_counts, np.int32)
    Q, qcap, _ = queries.shape
    K, kcap, _ = db.shape
    out = np.zeros((Q, K), np.int32)
    lib().orc_db_scores(_p(queries), Q, _p(qc), qcap, _p(db), K, _p(kc), kcap, _p(out), nthreads)
    return out


def db_topk(scores, k, id_base=0):
    scores = np.ascontiguousarray(scores, np.int32)
    Q, K = scores.shape
    os_ = np.zeros((Q, k), np.int32)
    oi = np.zeros((Q, k), np.int32)
    lib().orc_db_topk(_p(scores), Q, K, id_base, k, _p(os_), _p(oi))
    return os_, oi


def match_ratio_batch(frm, nfrom, to, nto, nthreads=1):
    frm = np.ascontiguousarray(frm, np.uint8)
    to = np.ascontiguousarray(to, np.uint8)
    P, cap, _ = frm.shape
    assert to.shape == frm.shape
    nf = np.ascontiguousarray(nfrom, np.int32)
    nt = np.ascontiguousarray(nto, np.int32)
    fi = np.zeros((P, cap), np.int64)
    ti = np.zeros((P, cap), np.int64)
    di = np.zeros((P, cap), np.int32)
    cnt = np.zeros(P, np.int32)
    lib().orc_match_ratio_batch(_p(frm), _p(nf), _p(to), _p(nt), P, cap, _p(fi), _p(ti), _p(di), _p(cnt), nthreads)
    return fi, ti, di, cnt


# TUM intrinsics the reference's file provider hard-codes (rgbd_file_provider.cpp:140-142)
TUM_CAMERA = dict(cx=319.5, cy=239.5, fx=525.0, fy=525.0, factor=np.float32(1.0) / np.float32(5000.0))


def backproject(xy, depth, cx, cy, fx, fy, factor):
    """pointsFromRgbdKeypoints: xy [n, 2] float64 (any row stride), depth [H, W] uint16 -> (xyz [n, 3] f64, valid [n] u8)"""
    xy = np.asarray(xy, np.float64)
    assert xy.ndim == 2 and xy.shape[1] == 2 and (xy.shape[0] == 0 or xy.strides[1] == 8)
    depth = np.ascontiguousarray(depth, np.uint16)
    n = xy.shape[0]
    xyz = np.zeros((max(n, 1), 3), np.float64)
    valid = np.zeros(max(n, 1), np.uint8)
    if n:
        lib().orc_backproject(C.c_void_p(xy.ctypes.data), n, xy.strides[0], _p(depth), depth.shape[1], depth.shape[0],
                              C.c_double(cx), C.c_double(cy), C.c_double(fx), C.c_double(fy), C.c_float(factor), _p(xyz),
                              _p(valid))
    return xyz[:n], valid[:n]


# ---- DistributedOrbOpenCvDetector (distributed_cv_feature.cpp), SURVEY 8(f) f1: dist_orb_oracle.c -----------------------
class DistParams(C.Structure):
    _fields_ = [("nlevels", C.c_int), ("ini_fast_thr", C.c_int), ("min_fast_thr", C.c_int), ("min_size", C.c_int),
                ("scale_factor", C.c_float)]


class DistDumps(C.Structure):
    _fields_ = [("pyr", C.c_void_p), ("blurred", C.c_void_p), ("n_fast", C.c_void_p), ("n_tree", C.c_void_p),
                ("passes", C.c_void_p)]


def dist_default_params(nlevels=8, ini_fast_thr=20, min_fast_thr=7, min_size=1000, scale_factor=1.2):
    """orb_params("orb", 1.2f, 8, 20, 7) + OrbExtractorPimpl(params, 1000)  (distributed_cv_feature.cpp:1184-1186)"""
    return DistParams(nlevels, ini_fast_thr, min_fast_thr, min_size, scale_factor)


def dist_level_geometry(W, H, nlevels=8, scale_factor=1.2):
    w = np.zeros(nlevels, np.int32)
    h = np.zeros(nlevels, np.int32)
    s = np.zeros(nlevels, np.float32)
    lib().orc_dist_level_geometry(W, H, nlevels, C.c_float(scale_factor), _p(w), _p(h), _p(s))
    return w, h, s


def resize_linear(img, dw, dh):
    img = np.ascontiguousarray(img, np.uint8)
    out = np.empty((dh, dw), np.uint8)
    lib().orc_resize_linear(_p(img), img.shape[1], img.shape[0], _p(out), dw, dh)
    return out


def fast_subimage(img, thr):
    """cv::FAST(img, thr, nonmax=True) -> [n, 3] (x, y, score), raster order"""
    img = np.ascontiguousarray(img, np.uint8)
    cap = img.size // 4 + 16
    out = np.empty((cap, 3), np.int32)
    n = lib().orc_fast_subimage(_p(img), img.shape[1], img.shape[1], img.shape[0], thr, _p(out), cap)
    return out[:n].copy()


def dist_fast_cells(img, ini_thr=20, min_thr=7):
    img = np.ascontiguousarray(img, np.uint8)
    cap = img.size // 4 + 64
    out = np.empty((cap, 3), np.int32)
    n = lib().orc_dist_fast_cells(_p(img), img.shape[1], img.shape[0], ini_thr, min_thr, _p(out), cap)
    return out[:n].copy()


def dist_tree(xys, w, h, scale_factor, min_size=1000):
    """distribute_keypoints_via_tree on a level of size (w, h) -> (kept indices in list order, passes)"""
    xys = np.ascontiguousarray(xys, np.int32)
    n = len(xys)
    out = np.empty(max(n, 1), np.int32)
    passes = C.c_int(0)
    if n == 0:
        return out[:0], 0
    m = lib().orc_dist_tree(_p(xys), n, 19, w - 19, 19, h - 19, C.c_float(scale_factor), C.c_uint(min_size), _p(out),
                            C.byref(passes))
    return out[:m].copy(), passes.value


def gauss7_u8(img):
    img = np.ascontiguousarray(img, np.uint8)
    out = np.empty_like(img)
    lib().orc_gauss7_u8(_p(img), img.shape[1], img.shape[0], _p(out))
    return out


def dist_cos(v):
    return np.float32(lib().orc_dist_cos(C.c_float(v)))


def dist_sin(v):
    return np.float32(lib().orc_dist_sin(C.c_float(v)))


def dist_describe(blurred, cx, cy, angle_deg):
    blurred = np.ascontiguousarray(blurred, np.uint8)
    out = np.empty(32, np.uint8)
    lib().orc_dist_describe(_p(blurred), blurred.shape[1], int(cx), int(cy), C.c_float(angle_deg), _p(out))
    return out


def dist_umax():
    u = np.zeros(16, np.int32)
    lib().orc_dist_umax(_p(u))
    return u


def dist_detect_gray(gray_img, params=None, cap=16384, dumps=False):
    """-> (keypoints[KP_DTYPE], descriptors[N,32]) in the reference's order (level, tree list order)"""
    g = np.ascontiguousarray(gray_img, np.uint8)
    H, W = g.shape
    prm = params or dist_default_params()
    kps = np.zeros(cap, KP_DTYPE)
    desc = np.zeros((cap, 32), np.uint8)
    d = None
    if dumps:
        w, h, _ = dist_level_geometry(W, H, prm.nlevels, prm.scale_factor)
        tot = int((w.astype(np.int64) * h).sum())
        pyr = np.zeros(tot, np.uint8)
        blr = np.zeros(tot, np.uint8)
        nf = np.zeros(prm.nlevels, np.int32)
        nt = np.zeros(prm.nlevels, np.int32)
        ps = np.zeros(prm.nlevels, np.int32)
        d = DistDumps(_p(pyr).value, _p(blr).value, _p(nf).value, _p(nt).value, _p(ps).value)
    n = lib().orc_dist_detect_gray(_p(g), W, H, C.byref(prm), _p(kps), _p(desc), cap, C.byref(d) if d else None)
    if n > cap:
        return dist_detect_gray(g, prm, cap=n, dumps=dumps)
    if not dumps:
        return kps[:n].copy(), desc[:n].copy()
    offs = np.concatenate([[0], np.cumsum(w.astype(np.int64) * h)])
    out = {
        "pyr": [pyr[offs[l]:offs[l + 1]].reshape(h[l], w[l]) for l in range(prm.nlevels)],
        "blurred": [blr[offs[l]:offs[l + 1]].reshape(h[l], w[l]) for l in range(prm.nlevels)],
        "n_fast": nf, "n_tree": nt, "passes": ps,
    }
    return kps[:n].copy(), desc[:n].copy(), out


def dist_detect_rgb(rgb, params=None, cap=16384):
    """DistributedOrbOpenCvDetector::detect restated (toGrayScale in memory order, then extract)"""
    return dist_detect_gray(gray(rgb), params, cap)


def dist_detect_rgb_batch(frames, params=None, cap=16384, nthreads=1):
    frames = np.ascontiguousarray(frames, np.uint8)
    B, H, W, _ = frames.shape
    prm = params or dist_default_params()
    kps = np.zeros((B, cap), KP_DTYPE)
    desc = np.zeros((B, cap, 32), np.uint8)
    counts = np.zeros(B, np.int32)
    lib().orc_dist_detect_rgb_batch(_p(frames), B, W, H, C.byref(prm), _p(kps), _p(desc), cap, _p(counts), nthreads)
    assert (counts <= cap).all(), "oracle batch capacity exceeded"
    return kps, desc, counts
