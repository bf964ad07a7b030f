"""Batched RANSAC PnP timing on the B200 box: frames/s of mslam_b200_pnp_ransac_batch_dev over 64 frames x
1000 correspondences (50 % outliers, 100 hypotheses), CUDA events on the context's stream, plus cv2.solvePnPRansac and the
numpy oracle on the host for the same scenes.  One JSON line -> stdout."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402


def main():
    pkg = graft.load_package()
    F, n, steps, warm = 64, 1000, 50, 5
    fx, fy, cx, cy = pkg.synth.TUM_K
    scenes = [pkg.synth.pnp_scene(1000 + i, n, 0.5, 0.7) for i in range(F)]
    obj = np.stack([s[0] for s in scenes])
    img = np.stack([s[1] for s in scenes])
    dev = torch.device("cuda:0")
    d_obj, d_img = torch.from_numpy(obj).to(dev), torch.from_numpy(img).to(dev)
    d_cnt = torch.full((F,), n, dtype=torch.int32, device=dev)
    d_rt = torch.zeros((F, 6), dtype=torch.float64, device=dev)
    d_pose = torch.zeros((F, 7), dtype=torch.float64, device=dev)
    d_inl = torch.zeros((F, n), dtype=torch.uint8, device=dev)
    d_info = torch.zeros((F, 4), dtype=torch.int32, device=dev)
    params = pkg.pnp_params(fx, fy, cx, cy, seed=7)
    stream = torch.cuda.Stream()
    with pkg.Context(max_width=64, max_height=64, max_batch=1, nfeatures=100, stream=stream.cuda_stream) as ctx:
        def step():
            ctx.pnp_ransac_batch_dev(d_obj, d_img, d_cnt, F, n, params, d_rt, d_pose, d_inl, d_info)
        with torch.cuda.stream(stream):
            for _ in range(warm):
                step()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record(stream)
            for _ in range(steps):
                step()
            e1.record(stream)
            torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        # end to end through the host entry point (host buffers, copies inside)
        t0 = time.perf_counter()
        for _ in range(10):
            rt, pose, inl, info = ctx.pnp_ransac_batch(obj, img, np.full(F, n, np.int32), params)
        e2e_ms = (time.perf_counter() - t0) / 10 * 1e3
    info = d_info.cpu().numpy()
    rt = d_rt.cpu().numpy()
    err = max(max(np.abs(rt[i, :3] - scenes[i][2]).max(), np.abs(rt[i, 3:] - scenes[i][3]).max()) for i in range(F))
    out = {"what": "batched RANSAC PnP, 64 frames x 1000 correspondences, 50 % outliers, 100 hypotheses + LM", "device_ms_per_batch": ms,
           "device_frames_per_s": F / ms * 1e3, "e2e_ms_per_batch": e2e_ms, "e2e_frames_per_s": F / e2e_ms * 1e3,
           "found": int(info[:, 0].sum()), "mean_inliers": float(info[:, 1].mean()), "mean_lm_iterations": float(info[:, 3].mean()),
           "max_abs_pose_error_vs_truth": float(err)}
    try:
        import cv2
        K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float32)
        cv2.setNumThreads(0)
        t0 = time.perf_counter()
        for i in range(16):
            r0, t0v = (scenes[i][2] + 0.02).reshape(3, 1).copy(), (scenes[i][3] + 0.05).reshape(3, 1).copy()
            cv2.solvePnPRansac(obj[i], img[i], K, None, r0, t0v, True, 100, 5.0, 0.99)
        out["cv2_frames_per_s_1_thread"] = 16 / (time.perf_counter() - t0)
        out["cv2_version"] = cv2.__version__
    except Exception as e:  # pragma: no cover
        out["cv2"] = repr(e)
    from oracle import pnp_oracle as PO
    t0 = time.perf_counter()
    for i in range(4):
        PO.solve_pnp_ransac(obj[i], img[i], fx, fy, cx, cy, seed=7, frame=i)
    out["oracle_frames_per_s"] = 4 / (time.perf_counter() - t0)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
