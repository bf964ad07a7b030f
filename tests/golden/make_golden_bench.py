#!/usr/bin/env python
"""Golden fixtures on the inputs bench.py itself runs: OpenCV (cv2, the library behind the reference's
OrbOpenCvDetector / OrbOpenCvMatcher -- orb_feature.cpp:25,40,81,96) on frames of modular-slam_b200/synth.py's
generator (seed 1234, the bench generator) at the three configured shapes

    640x480 / 1000 keypoints (configs[1]), 1280x720 / 2000 (configs[2]), 1920x1080 / 4000 (configs[3])

two consecutive frames (t = 0, 1) of sequence 0 each: keypoints (all six fields), descriptors, kNN k=2, the
reference's ratio-0.7 matches and the cross-checked matches.  The synthetic textures have ~12x the FAST corner density
of a natural frame, so retainBest() cuts through ties on every pyramid level -- the regime the bundled-frame goldens
(make_golden.py) do not reach.

The frames themselves are NOT stored: synth.frame_batch is seeded numpy, the fixture keeps a sha256 of each frame so
that a change of the generator is noticed.  Gray conversion = the numpy statement of frame.cpp:18-19, not the oracle.

Needs cv2 (version recorded in the file; 4.13.0 here, the reference pins 4.8.1) -> run in the build container only;
tests/golden/bench_synth.npz is committed.
"""
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
from make_golden import cv_cross, cv_knn2, cv_match_ratio, cv_orb, ref_gray  # noqa: E402

SHAPES = ((640, 480, 1000), (1280, 720, 2000), (1920, 1080, 4000))


def main():
    cv2.setNumThreads(1)
    synth = graft.load_package().synth
    out = {"cv2_version": np.array(cv2.__version__)}
    for w, h, n in SHAPES:
        tag = f"{w}x{h}"
        descs = []
        bases = None
        for t in (0, 1):
            frames, bases = synth.frame_batch(w, h, 1, t, seed=1234, bases=bases)
            kp, d = cv_orb(ref_gray(frames[0]), n)
            out[f"{tag}_sha{t}"] = np.array(hashlib.sha256(frames[0].tobytes()).hexdigest())
            out[f"{tag}_kp{t}"], out[f"{tag}_desc{t}"] = kp, d
            descs.append(d)
            print(tag, t, len(kp), np.bincount(kp["octave"]).tolist())
        out[f"{tag}_ratio_0_to_1"] = cv_match_ratio(descs[0], descs[1])  # OrbOpenCvMatcher::match(from = t0, to = t1)
        out[f"{tag}_knn2_q1_t0"] = cv_knn2(descs[1], descs[0])
        out[f"{tag}_cross_q1_t0"] = cv_cross(descs[1], descs[0])
        print(tag, "ratio", len(out[f"{tag}_ratio_0_to_1"]), "cross", len(out[f"{tag}_cross_q1_t0"]))
    path = os.path.join(HERE, "bench_synth.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path))


if __name__ == "__main__":
    sys.exit(main())
