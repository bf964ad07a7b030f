#!/usr/bin/env python
"""Golden fixtures for cv::ORB with parameters OTHER than the reference's hard-coded ones (orb_feature.cpp:25 uses
cv::ORB::create(1000) only, but the C ABI's mslam_b200_config exposes nlevels / scale_factor / fast_threshold /
edge_threshold / nfeatures): cv2.ORB_create(...) on synth.frame_batch frames, keypoints (six fields) + descriptors.
Includes scale factors of 2 and above, where the pyramid's quad-table resize kernel hands over to its per-pixel fallback.
Frames are not stored (seeded numpy); a sha256 of each guards the generator.  Needs cv2 -> run in the build container only;
tests/golden/orb_params.npz is committed."""
import hashlib
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
import __graft_entry__ as graft  # noqa: E402
from make_golden import KP, ref_gray  # noqa: E402

# (nlevels, scale_factor, fast_threshold, edge_threshold, nfeatures, W, H)
CASES = ((5, 1.5, 30, 31, 700, 640, 480), (3, 2.5, 20, 31, 500, 900, 700), (11, 1.1, 15, 31, 1500, 512, 384),
         (4, 2.0, 25, 40, 300, 800, 600), (8, 1.2, 20, 31, 1000, 333, 500), (2, 3.0, 10, 31, 400, 640, 480))


def main():
    cv2.setNumThreads(1)
    synth = graft.load_package().synth
    out = {"cv2_version": np.array(cv2.__version__), "cases": np.array(CASES, np.float64)}
    for i, (nl, sf, thr, edge, nf, w, h) in enumerate(CASES):
        frames, _ = synth.frame_batch(w, h, 1, 0, seed=10 * nl + thr)
        orb = cv2.ORB_create(nfeatures=nf, scaleFactor=sf, nlevels=nl, edgeThreshold=edge, fastThreshold=thr)
        k, d = orb.detectAndCompute(ref_gray(frames[0]), None)
        a = np.zeros(len(k), KP)
        for j, p in enumerate(k):
            a[j] = (p.pt[0], p.pt[1], p.size, p.angle, p.response, p.octave)
        out[f"c{i}_sha"] = np.array(hashlib.sha256(frames[0].tobytes()).hexdigest())
        out[f"c{i}_kp"], out[f"c{i}_desc"] = a, d
        print(i, (nl, sf, thr, edge, nf, w, h), len(a), np.bincount(a["octave"]).tolist())
    np.savez_compressed(os.path.join(HERE, "orb_params.npz"), **out)


if __name__ == "__main__":
    main()
