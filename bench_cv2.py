"""bench_cv2.py -- OpenCV itself (cv2, the library the reference's detector / matcher wrap: orb_feature.cpp:25,40,81,96) as an
informational CPU figure next to the C port: one frame pair per worker PROCESS, cv2 single-threaded inside each, all host
cores busy -- the fairest form of "the reference on this box's CPU" that can be run without the reference's C++ build.
Import-light on purpose (numpy + cv2 only): the workers are spawned."""
import time

import numpy as np


def _gray(bgr):
    f = bgr.astype(np.float32)  # frame.cpp:18-19
    return np.minimum(np.float32(255), (np.float32(0.299) * f[..., 0] + np.float32(0.587) * f[..., 1]) + np.float32(0.114) * f[..., 2]).astype(np.uint8)


def _work(args):
    import cv2
    prev, cur, nfeat, mode, reps = args
    cv2.setNumThreads(1)
    orb = cv2.ORB_create(nfeat)
    bf = cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=(mode == 1))
    _, d0 = orb.detectAndCompute(_gray(prev), None)
    n = 0
    t0 = time.perf_counter()
    for _ in range(reps):
        _, d1 = orb.detectAndCompute(_gray(cur), None)  # toGrayScale + detectAndCompute, as OrbOpenCvDetector::detect
        m = bf.match(d1, d0) if mode == 1 else bf.knnMatch(d1, d0, 2)
        n += len(m)
    return time.perf_counter() - t0, n


def pool_frames_per_s(frames_prev, frames_cur, nfeat, mode, procs, reps=3, timeout=120):
    """frames/s of `procs` worker processes (python bench_cv2.py ...), each running detect + match on its own frame pair `reps`
    times after a common start time; None if cv2 is absent or a worker fails"""
    try:
        import cv2  # noqa: F401
    except ImportError:
        return None
    import os
    import subprocess
    import sys
    import tempfile
    n = min(len(frames_cur), procs)
    with tempfile.TemporaryDirectory() as d:
        for i in range(n):
            np.save(os.path.join(d, f"p{i}.npy"), np.stack([frames_prev[i], frames_cur[i]]))
        go = time.time() + 6.0  # imports + warm-up are over by then on every worker
        ps = [subprocess.Popen([sys.executable, os.path.abspath(__file__), os.path.join(d, f"p{i}.npy"), str(nfeat), str(mode),
                                str(reps), repr(go)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True) for i in range(n)]
        worst = 0.0
        try:
            for p in ps:
                out, _ = p.communicate(timeout=timeout)
                if p.returncode != 0:
                    return None
                worst = max(worst, float(out.strip().split()[-1]))
        except (subprocess.TimeoutExpired, ValueError, IndexError):
            for p in ps:
                p.kill()
            return None
    return n * reps / worst if worst > 0 else None


if __name__ == "__main__":
    import sys
    pair = np.load(sys.argv[1])
    nfeat, mode, reps, go = int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), float(sys.argv[5])
    _work((pair[0], pair[1], nfeat, mode, 1))
    while time.time() < go:
        time.sleep(0.001)
    t0 = time.time()
    _work((pair[0], pair[1], nfeat, mode, reps))
    print(time.time() - go)  # from the COMMON start to this worker's end
