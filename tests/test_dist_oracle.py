"""CPU tests of oracle/dist_orb_oracle.c (the restatement of DistributedOrbOpenCvDetector):
every OpenCV primitive against the committed cv2 vectors (tests/golden/dist_primitives.npz, made by
tests/golden/make_golden_dist.py), and the quad-tree against a second, independent restatement written here with Python
lists used the way the reference uses std::list (distributed_cv_feature.cpp:941-1106).  Nothing reads /root/reference."""
import hashlib
import math
import os

import numpy as np
import pytest

from conftest import GOLDEN


@pytest.fixture(scope="module")
def prim():
    return np.load(os.path.join(GOLDEN, "dist_primitives.npz"))


def _sha(a):
    return hashlib.sha1(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_pattern_table_was_checked_against_the_reference(prim):
    assert bool(prim["pattern_checked_equal"])


def test_level_geometry(oracle, prim):
    w, h, s = oracle.dist_level_geometry(311, 237)
    assert (np.stack([w, h], 1) == prim["img_level_sizes"]).all()
    w, h, s = oracle.dist_level_geometry(640, 480)
    assert (np.stack([w, h], 1) == prim["bundled_level_sizes"]).all()
    # float product chain, not pow (orb_params::calc_scale_factors)
    f = np.float32(1.0)
    for l in range(8):
        assert s[l] == f
        f = np.float32(np.float32(1.2) * f)


def test_resize_linear_matches_cv2(oracle, prim, bundled):
    cur = prim["img"]
    for l in range(1, 8):
        ref = prim[f"img_level{l}"]
        got = oracle.resize_linear(cur, ref.shape[1], ref.shape[0])
        assert (got == ref).all(), f"level {l}: {(got != ref).sum()} px differ"
        cur = ref
    cur = oracle.gray(bundled["bgr0"])
    for l in range(1, 8):
        w, h = prim["bundled_level_sizes"][l]
        cur = oracle.resize_linear(cur, int(w), int(h))
        assert _sha(cur) == str(prim["bundled_level_sha1"][l - 1]), f"bundled level {l}"


def test_gauss7_matches_cv2(oracle, prim, bundled):
    assert (oracle.gauss7_u8(prim["img"]) == prim["img_gauss"]).all()
    cur = oracle.gray(bundled["bgr0"])
    for l in range(8):
        if l:
            w, h = prim["bundled_level_sizes"][l]
            cur = oracle.resize_linear(cur, int(w), int(h))
        assert _sha(oracle.gauss7_u8(cur)) == str(prim["bundled_gauss_sha1"][l]), f"bundled level {l}"


@pytest.mark.parametrize("thr", [20, 7])
def test_fast_on_subimages_matches_cv2(oracle, prim, thr):
    img = prim["img"]
    xys, offs = prim[f"fast{thr}_xys"], prim[f"fast{thr}_offs"]
    for i, (y0, y1, x0, x1) in enumerate(prim["fast_rects"]):
        got = oracle.fast_subimage(np.ascontiguousarray(img[y0:y1, x0:x1]), thr)
        ref = xys[offs[i]:offs[i + 1]]
        assert got.shape == ref.shape and (got == ref).all(), f"rect {i} thr {thr}"


def test_cell_loop_matches_cv2_per_cell(oracle, prim, bundled):
    """compute_fast_keypoints' cell cutting + threshold fallback: oracle vs cv2.FAST run cell by cell by the generator"""
    cur = oracle.gray(bundled["bgr0"])
    for l in range(4):
        if l:
            w, h = prim["bundled_level_sizes"][l]
            cur = oracle.resize_linear(cur, int(w), int(h))
        if l in (0, 3):
            got = oracle.dist_fast_cells(cur)
            ref = prim[f"bundled_cells_level{l}"]
            assert got.shape == ref.shape and (got == ref).all(), f"level {l}"


def test_umax_is_the_orb_table(oracle):
    assert list(oracle.dist_umax()) == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]


def test_polynomial_cos_sin(oracle):
    f = np.float32
    for v in np.linspace(-1.0, 7.0, 801, dtype=np.float32):
        c, s = oracle.dist_cos(v), oracle.dist_sin(v)
        assert abs(float(c) - math.cos(float(v))) < 2e-3 and abs(float(s) - math.sin(float(v))) < 2e-3
    # restated in numpy f32, operation by operation (util::cos, :465-496)
    PI = f(3.14159265358979)
    PI_2, TWO_PI = f(PI / f(2.0)), f(f(2.0) * PI)
    INV, T32 = f(f(1.0) / TWO_PI), f(f(3.0) * PI_2)

    def pc(v):
        v2 = f(v * v)
        return f(f(0.99940307) + f(v2 * f(f(-0.49558072) + f(f(0.03679168) * v2))))

    def cos32(v):
        v = f(v - f(f(math.floor(f(v * INV))) * TWO_PI))
        v = v if f(0.0) < v else f(-v)
        if v < PI_2:
            return pc(v)
        if v < PI:
            return f(-pc(f(PI - v)))
        if v < T32:
            return f(-pc(f(v - PI)))
        return pc(f(TWO_PI - v))
    rng = np.random.default_rng(0)
    for v in rng.uniform(0, 6.2832, 2000).astype(np.float32):
        assert oracle.dist_cos(v) == cos32(v)
        assert oracle.dist_sin(v) == cos32(f(PI_2 - v))


# ---- the tree, restated a second time with Python lists as std::list ---------------------------------------------------
def _py_tree(xys, w, h, sf, min_size=1000):
    f = np.float32
    min_x = min_y = 19
    max_x, max_y = w - 19, h - 19
    ratio = (max_x - min_x) / (max_y - min_y)
    if ratio > 1:
        nx, ny = int(math.floor(ratio + 0.5)), 1
        dx, dy = (max_x - min_x) / nx, float(max_y - min_y)
    else:
        nx, ny = 1, int(math.floor(1 / ratio + 0.5))
        dx, dy = float(max_x - min_y), (max_y - min_y) / ny
    nodes = []
    for i in range(nx * ny):
        ix, iy = i % nx, i // nx
        nodes.append({"b": (int(dx * ix), int(dy * iy)), "e": (int(dx * (ix + 1)), int(dy * (iy + 1))), "k": []})
    for k, (x, y, _) in enumerate(xys):
        nodes[int(float(f(x)) / dx) + int(float(f(y)) / dy) * nx]["k"].append(k)
    nodes = [n for n in nodes if n["k"]]
    passes = 0
    while True:
        prev = len(nodes)
        i = 0
        while i < len(nodes):
            n = nodes[i]
            area = (n["e"][0] - n["b"][0]) * (n["e"][1] - n["b"][1])
            if len(n["k"]) == 1 or f(f(f(area) * f(sf)) * f(sf)) <= f(min_size):
                i += 1
                continue
            hx = math.ceil((n["e"][0] - n["b"][0]) / 2.0)
            hy = math.ceil((n["e"][1] - n["b"][1]) / 2.0)
            (bx, by), (ex, ey) = n["b"], n["e"]
            ch = [{"b": (bx, by), "e": (bx + hx, by + hy), "k": []}, {"b": (bx + hx, by), "e": (ex, by + hy), "k": []},
                  {"b": (bx, by + hy), "e": (bx + hx, ey), "k": []}, {"b": (bx + hx, by + hy), "e": (ex, ey), "k": []}]
            for k in n["k"]:
                ch[(1 if bx + hx <= xys[k][0] else 0) + (2 if by + hy <= xys[k][1] else 0)]["k"].append(k)
            del nodes[i]          # erase: the iterator moves to the successor ...
            pushed = 0
            for c in ch:
                if c["k"]:
                    nodes.insert(0, c)  # ... and the children go to the front
                    pushed += 1
            i += pushed
        passes += 1
        if len(nodes) == prev:
            break
    keep = []
    for n in nodes:
        best = n["k"][0]
        for k in n["k"][1:]:
            if xys[k][2] > xys[best][2]:
                best = k
        keep.append(best)
    return keep, passes


@pytest.mark.parametrize("w,h,sf,n,seed", [(640, 480, 1.0, 1500, 0), (533, 400, 1.2, 900, 1), (179, 134, 3.583181, 200, 2),
                                           (480, 640, 1.0, 1200, 3), (1280, 720, 1.0, 3000, 4), (311, 237, 1.44, 40, 5),
                                           (640, 480, 1.0, 1, 6), (640, 480, 1.0, 2, 7)])
def test_tree_against_a_second_restatement(oracle, w, h, sf, n, seed):
    rng = np.random.default_rng(seed)
    # distinct integer positions inside cv::FAST's reach, relative to the border origin; clustered so that deep
    # subdivisions, single-child divisions and response ties all occur
    pts = set()
    while len(pts) < n:
        if rng.random() < 0.5:
            x, y = rng.integers(3, w - 41), rng.integers(3, h - 41)
        else:
            cx, cy = rng.integers(3, w - 41), rng.integers(3, h - 41)
            x = int(np.clip(cx + rng.integers(-6, 7), 3, w - 42))
            y = int(np.clip(cy + rng.integers(-6, 7), 3, h - 42))
        pts.add((int(x), int(y)))
    pts = sorted(pts, key=lambda p: ((p[1] - 3) // 64, (p[0] - 3) // 64, p[1], p[0]))  # the cell loop's arrival order
    xys = np.array([(x, y, rng.integers(7, 40)) for x, y in pts], np.int32)
    keep, passes = oracle.dist_tree(xys, w, h, np.float32(sf))
    ref_keep, ref_passes = _py_tree([tuple(int(v) for v in r) for r in xys], w, h, np.float32(sf))
    assert list(keep) == ref_keep
    assert passes == ref_passes
    assert len(set(keep)) == len(keep)


def test_whole_detector_runs_and_is_deterministic(oracle, bundled):
    kp, desc, d = oracle.dist_detect_gray(oracle.gray(bundled["bgr0"]), dumps=True)
    kp2, desc2 = oracle.dist_detect_rgb(bundled["bgr0"])
    assert kp.tobytes() == kp2.tobytes() and (desc == desc2).all()
    assert len(kp) == int(d["n_tree"].sum()) and (d["n_tree"] <= d["n_fast"]).all()
    assert (np.diff(kp["octave"]) >= 0).all()  # level-major
    # every keypoint sits inside cv::FAST's reach of its level and carries that level's patch size
    w, h, s = oracle.dist_level_geometry(640, 480)
    for l in range(8):
        m = kp["octave"] == l
        x, y = kp["x"][m] / (s[l] if l else 1), kp["y"][m] / (s[l] if l else 1)
        assert (x > 21.5).all() and (x < w[l] - 21.5).all() and (y > 21.5).all() and (y < h[l] - 21.5).all()
        assert (kp["size"][m] == np.float32(int(31 * s[l]))).all()
