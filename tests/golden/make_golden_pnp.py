"""Golden outputs of cv2.solvePnPRansac (opencv-python-headless 4.13.0; the reference pins OpenCV 4.8.1) for the call
OpenCvRansacPnp::solvePnp makes (src/lib/modular_slam/cv_ransac_pnp.cpp:56-57):
    cv::solvePnPRansac(obj f32, img f32, K f32, noArray(), rvec, tvec, useExtrinsicGuess=true, 100, 5.0, 0.99, inliers)
on synthetic scenes from modular-slam_b200/synth.py::pnp_scene.  Run HERE (needs cv2):  python tests/golden/make_golden_pnp.py
-> tests/golden/pnp.npz"""
import os
import sys

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402

CASES = [  # (seed, n, outlier fraction, pixel noise)
    (0, 300, 0.3, 0.5), (1, 300, 0.3, 0.5), (2, 300, 0.3, 0.5), (3, 300, 0.3, 0.5),
    (10, 1000, 0.5, 0.7), (11, 1000, 0.1, 0.3), (12, 693, 0.4, 0.5),  # 693 = matches of the bundled frame pair
    (20, 40, 0.2, 0.3), (21, 24, 0.25, 0.2), (22, 100, 0.0, 0.5),
]


def main():
    synth = graft.load_package().synth
    fx, fy, cx, cy = synth.TUM_K
    K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float32)
    out = {"cases": np.array(CASES, np.float64), "cv2_version": np.array(cv2.__version__)}
    for i, (seed, n, of, noise) in enumerate(CASES):
        X, x, rv, tv, _ = synth.pnp_scene(seed, n, of, noise)
        # the extrinsic guess the frontend passes is the previous pose: truth plus a small offset
        r0, t0 = (rv + 0.02).reshape(3, 1).copy(), (tv + 0.05).reshape(3, 1).copy()
        ok, r, t, inl = cv2.solvePnPRansac(X, x, K, None, r0, t0, True, 100, 5.0, 0.99)
        assert ok
        mask = np.zeros(n, np.uint8)
        mask[inl.ravel()] = 1
        out[f"obj{i}"], out[f"img{i}"] = X, x
        out[f"rvec{i}"], out[f"tvec{i}"], out[f"inl{i}"] = r.ravel(), t.ravel(), mask
        out[f"true_rvec{i}"], out[f"true_tvec{i}"] = rv, tv
        print(i, n, int(mask.sum()), np.abs(r.ravel() - rv).max(), np.abs(t.ravel() - tv).max())
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "pnp.npz"), **out)


if __name__ == "__main__":
    main()
