#!/usr/bin/env python
"""Generate the golden fixtures in tests/golden/ by running OpenCV itself (cv2, the library the reference's
OrbOpenCvDetector / OrbOpenCvMatcher wrap -- orb_feature.cpp:25,40,81,96) in the build container.

Inputs : the reference's bundled frames /root/reference/data/rgb/000{0,1}.png (identical copies under test/data/),
         read exactly as RgbdFileProvider does (cv::imread -> BGR bytes, rgbd_file_provider.cpp:66).
Outputs: tests/golden/bundled.npz   (frames + cv2 keypoints / descriptors / matches)
         tests/golden/synthetic.npz (seeded odd-size noise frame; tie-heavy descriptor sets with cv2 answers)
Gray conversion uses the numpy statement of frame.cpp:18-19 (three rounded f32 products, two rounded adds,
min(255), truncate), NOT the oracle, so the fixtures do not depend on the code they test.

Needs /root/reference and cv2 (recorded version below) -> run here only; the .npz files are committed.
"""
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
KP = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4")])


def ref_gray(bgr):
    f = bgr.astype(np.float32)
    v = (np.float32(0.299) * f[..., 0] + np.float32(0.587) * f[..., 1]) + np.float32(0.114) * f[..., 2]
    return np.minimum(np.float32(255), v).astype(np.uint8)


def cv_orb(gray, n):
    k, d = cv2.ORB_create(n).detectAndCompute(gray, None)
    a = np.zeros(len(k), KP)
    for i, p in enumerate(k):
        a[i] = (p.pt[0], p.pt[1], p.size, p.angle, p.response, p.octave)
    return a, d  # native cv2 order


def cv_match_ratio(frm, to):
    knn = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(to, frm, 2)
    return np.array([(m[0].trainIdx, m[0].queryIdx, int(m[0].distance)) for m in knn
                     if m[0].distance < 0.7 * m[1].distance], np.int32).reshape(-1, 3)


def cv_knn2(q, t):
    knn = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(q, t, 2)
    return np.array([(m[0].trainIdx, int(m[0].distance), m[1].trainIdx, int(m[1].distance)) for m in knn], np.int32)


def cv_cross(q, t):
    m = cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=True).match(q, t)
    return np.array([(x.queryIdx, x.trainIdx, int(x.distance)) for x in m], np.int32).reshape(-1, 3)


def main():
    cv2.setNumThreads(1)
    out = {"cv2_version": np.array(cv2.__version__)}
    descs = []
    for i, name in enumerate(["0000", "0001"]):
        bgr = cv2.imread(f"/root/reference/data/rgb/{name}.png")
        g = ref_gray(bgr)
        kp, d = cv_orb(g, 1000)
        out[f"bgr{i}"] = bgr
        out[f"kp{i}"] = kp
        out[f"desc{i}"] = d
        descs.append(d)
    out["match_0_to_1"] = cv_match_ratio(descs[0], descs[1])      # OrbOpenCvMatcher::match(from=frame0, to=frame1)
    out["cross_q1_t0"] = cv_cross(descs[1], descs[0])
    out["knn2_q1_t0"] = cv_knn2(descs[1], descs[0])
    np.savez_compressed(os.path.join(HERE, "bundled.npz"), **out)
    # the bundled depth frames (data/depth, 16-bit PNG as the reference's RgbdFileProvider reads them with IMREAD_UNCHANGED,
    # rgbd_file_provider.cpp:95) for the back-projection step; kept in a separate file so bundled.npz stays untouched
    np.savez_compressed(os.path.join(HERE, "depth.npz"),
                        depth0=cv2.imread("/root/reference/data/depth/0000.png", cv2.IMREAD_UNCHANGED),
                        depth1=cv2.imread("/root/reference/data/depth/0001.png", cv2.IMREAD_UNCHANGED))

    rng = np.random.default_rng(20261104)
    syn = {"cv2_version": np.array(cv2.__version__)}
    noise = cv2.GaussianBlur(rng.integers(0, 256, (237, 311), dtype=np.uint8), (5, 5), 1.0)
    kp, d = cv_orb(noise, 300)
    syn["noise_gray"], syn["noise_kp"], syn["noise_desc"] = noise, kp, d
    # every stage primitive that cv2 exposes, on the same frame
    syn["noise_fast"] = np.array([(int(p.pt[0]), int(p.pt[1]), int(p.response)) for p in
                                  cv2.FastFeatureDetector_create(20, True).detect(noise)], np.int32)
    syn["noise_resize"] = cv2.resize(noise, (259, 198), interpolation=cv2.INTER_LINEAR_EXACT)
    k1 = cv2.getGaussianKernel(7, 2, cv2.CV_32F)
    syn["noise_blur"] = cv2.sepFilter2D(noise, cv2.CV_8U, k1, k1, borderType=cv2.BORDER_REFLECT_101)
    a = rng.integers(0, 256, (500, 32), dtype=np.uint8)
    b = rng.integers(0, 256, (300, 32), dtype=np.uint8)
    a[:, 4:] = 0
    b[:, 4:] = 0  # only 32 bits vary -> many distance ties
    syn["tie_a"], syn["tie_b"] = a, b
    syn["tie_knn2"] = cv_knn2(a, b)
    syn["tie_cross"] = cv_cross(a, b)
    syn["tie_ratio"] = cv_match_ratio(b, a)
    np.savez_compressed(os.path.join(HERE, "synthetic.npz"), **syn)
    for f in ("bundled.npz", "synthetic.npz"):
        print(f, os.path.getsize(os.path.join(HERE, f)))


if __name__ == "__main__":
    sys.exit(main())
