#!/usr/bin/env python
"""Generate tests/golden/dist_primitives.npz: outputs of OpenCV itself (cv2) for every OpenCV primitive that the
reference's DistributedOrbOpenCvDetector calls (src/lib/modular_slam/distributed_cv_feature.cpp):
    cv::resize(..., INTER_LINEAR)            :838
    cv::FAST(cell, thr, true), thr 20 and 7   :915-922
    cv::GaussianBlur(7x7, 2, 2, REFLECT_101)  :797-798   (u8 fixed-point path)
The detector as a whole has no cv2 counterpart (its control logic is in the reference tree and the reference cannot be
built here), so these primitive vectors are what pins oracle/dist_orb_oracle.c.

Inputs : a seeded 311x237 blurred-noise image (stored in the fixture) and the reference's bundled frame
         /root/reference/data/rgb/0000.png (only SHA-1 digests of the cv2 outputs are stored for it; the frame itself is
         already in tests/golden/bundled.npz).
Also checks, at generation time, that the reference's orb_point_pairs table (:25-282) is the same 1024 numbers as
modular-slam_b200/csrc/brief_pattern.inc.   Needs /root/reference and cv2 -> run here only; the .npz is committed.
"""
import hashlib
import os
import re
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))


def ref_gray(bgr):
    f = bgr.astype(np.float32)
    v = (np.float32(0.299) * f[..., 0] + np.float32(0.587) * f[..., 1]) + np.float32(0.114) * f[..., 2]
    return np.minimum(np.float32(255), v).astype(np.uint8)


def level_sizes(W, H, nlevels=8, sf=np.float32(1.2)):
    """orb_params::calc_scale_factors (:413-421) + compute_image_pyramid (:835-836), restated in numpy"""
    s = np.float32(1.0)
    out = [(W, H)]
    for _ in range(1, nlevels):
        s = np.float32(sf * s)
        sc = float(s)

        def rnd(v):  # std::round: half away from zero
            return int(np.floor(v + 0.5)) if v >= 0 else -int(np.floor(-v + 0.5))
        out.append((rnd(W * 1.0 / sc), rnd(H * 1.0 / sc)))
    return out


def fast_cells(img, rects, thr):
    fd = cv2.FastFeatureDetector_create(threshold=thr, nonmaxSuppression=True)
    rows, offs = [], [0]
    for (y0, y1, x0, x1) in rects:
        k = fd.detect(img[y0:y1, x0:x1])  # a view with the parent's stride, like rowRange().colRange()
        rows += [(int(p.pt[0]), int(p.pt[1]), int(p.response)) for p in k]
        offs.append(len(rows))
    return np.array(rows, np.int32).reshape(-1, 3), np.array(offs, np.int32)


def sha(a):
    return hashlib.sha1(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    cv2.setNumThreads(1)
    out = {"cv2_version": np.array(cv2.__version__)}
    # the pattern table
    src = open("/root/reference/src/lib/modular_slam/distributed_cv_feature.cpp").read()
    i = src.index("orb_point_pairs[orb_point_pairs_size]")
    body = re.sub(r"/\*.*?\*/", "", src[src.index("{", i) + 1:src.index("};", i)], flags=re.S)
    theirs = [int(v) for v in re.findall(r"-?\d+", body)]
    t = open(os.path.join(ROOT, "modular-slam_b200", "csrc", "brief_pattern.inc")).read()
    t = re.sub(r"//.*", "", re.sub(r"/\*.*?\*/", "", t, flags=re.S))
    mine = [int(v) for v in re.findall(r"-?\d+", t)]
    assert len(theirs) == 1024 and theirs == mine, "orb_point_pairs differs from brief_pattern.inc"
    out["pattern_checked_equal"] = np.array(True)

    rng = np.random.default_rng(3)
    img = cv2.GaussianBlur(rng.integers(0, 256, (237, 311), dtype=np.uint8), (5, 5), 1.2)
    out["img"] = img
    sizes = level_sizes(311, 237)
    out["img_level_sizes"] = np.array(sizes, np.int32)
    cur = img
    for l in range(1, 8):
        cur = cv2.resize(cur, sizes[l], interpolation=cv2.INTER_LINEAR)
        out[f"img_level{l}"] = cur
    out["img_gauss"] = cv2.GaussianBlur(img.copy(), (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
    rects = [(19, 89, 19, 89), (83, 153, 147, 217), (19, 60, 19, 26), (100, 170, 200, 241), (147, 218, 275, 292),
             (0, 237, 0, 311)]
    out["fast_rects"] = np.array(rects, np.int32)
    for thr in (20, 7):
        k, o = fast_cells(img, rects, thr)
        out[f"fast{thr}_xys"], out[f"fast{thr}_offs"] = k, o

    # bundled frame: digests only
    g = ref_gray(cv2.imread("/root/reference/data/rgb/0000.png"))
    sizes = level_sizes(640, 480)
    out["bundled_level_sizes"] = np.array(sizes, np.int32)
    dig = []
    cur = g
    levels = [g]
    for l in range(1, 8):
        cur = cv2.resize(cur, sizes[l], interpolation=cv2.INTER_LINEAR)
        levels.append(cur)
        dig.append(sha(cur))
    out["bundled_level_sha1"] = np.array(dig)
    out["bundled_gauss_sha1"] = np.array([sha(cv2.GaussianBlur(lv.copy(), (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101))
                                          for lv in levels])
    # cv2.FAST on every cell of level 0 and level 3 exactly as compute_fast_keypoints cuts them (:857-922), thr 20 -> 7
    for l in (0, 3):
        lv = levels[l]
        h, w = lv.shape
        cells = []
        for i in range((h - 38) // 64 + 1):
            y0 = 19 + 64 * i
            if h - 19 - 6 <= y0:
                continue
            y1 = min(y0 + 70, h - 19)
            for j in range((w - 38) // 64 + 1):
                x0 = 19 + 64 * j
                if w - 19 - 6 <= x0:
                    continue
                x1 = min(x0 + 70, w - 19)
                cells.append((y0, y1, x0, x1, i, j))
        rows = []
        for (y0, y1, x0, x1, i, j) in cells:
            k, _ = fast_cells(lv, [(y0, y1, x0, x1)], 20)
            if len(k) == 0:
                k, _ = fast_cells(lv, [(y0, y1, x0, x1)], 7)
            k = k.copy()
            k[:, 0] += 64 * j
            k[:, 1] += 64 * i
            rows.append(k)
        out[f"bundled_cells_level{l}"] = np.concatenate(rows).astype(np.int32)
    np.savez_compressed(os.path.join(HERE, "dist_primitives.npz"), **out)
    print("wrote dist_primitives.npz:", {k: (v.shape if hasattr(v, "shape") else v) for k, v in out.items()})


if __name__ == "__main__":
    sys.exit(main())
