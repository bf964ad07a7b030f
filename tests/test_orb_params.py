"""cv::ORB with parameters other than the reference's hard-coded cv::ORB::create(1000) (orb_feature.cpp:25): the fields of
mslam_b200_config -- nlevels, scale_factor (incl. 2 and above, where the quad-table resize kernel hands over to its
per-pixel fallback), fast_threshold, edge_threshold, nfeatures -- against OpenCV itself (tests/golden/orb_params.npz, made
by tests/golden/make_golden_params.py with cv2 4.13.0).  Bar: bit-exact keypoints (raw bytes, canonical order) and
descriptors.  CPU half: the oracle; GPU half: the CUDA path through the C ABI."""
import hashlib
import os

import numpy as np
import pytest

from conftest import GOLDEN, canon


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLDEN, "orb_params.npz"))


def _cases(gold):
    return [(i, int(c[0]), float(c[1]), int(c[2]), int(c[3]), int(c[4]), int(c[5]), int(c[6])) for i, c in enumerate(gold["cases"])]


def _frame(pkg, gold, i, nl, thr, w, h):
    f, _ = pkg.synth.frame_batch(w, h, 1, 0, seed=10 * nl + thr)
    assert hashlib.sha256(f[0].tobytes()).hexdigest() == str(gold[f"c{i}_sha"]), \
        "synth.frame_batch changed: regenerate tests/golden/orb_params.npz (make_golden_params.py)"
    return f[0]


def _same(kp, desc, gk, gd, what):
    assert len(kp) == len(gk), f"{what}: {len(kp)} keypoints, cv2 has {len(gk)}"
    if kp.tobytes() != gk.tobytes():
        bad = [j for j in range(len(kp)) if kp[j].tobytes() != gk[j].tobytes()]
        raise AssertionError(f"{what}: {len(bad)} keypoints differ from cv2, first {bad[0]}: {kp[bad[0]]} vs {gk[bad[0]]}")
    nbad = int((desc != gd).any(axis=1).sum())
    assert nbad == 0, f"{what}: {nbad}/{len(desc)} descriptors differ from cv2"


def test_oracle_equals_cv2_for_non_default_parameters(pkg, oracle, gold):
    for i, nl, sf, thr, edge, nf, w, h in _cases(gold):
        kp, desc = oracle.orb_detect_rgb(_frame(pkg, gold, i, nl, thr, w, h), oracle.default_params(nf, nl, sf, thr, edge))
        gk, gd = canon(gold[f"c{i}_kp"], gold[f"c{i}_desc"])
        _same(kp, desc, gk.astype(oracle.KP_DTYPE), gd, f"oracle, case {i}")


@pytest.mark.gpu
def test_cuda_equals_cv2_for_non_default_parameters(pkg, oracle, gold):
    for i, nl, sf, thr, edge, nf, w, h in _cases(gold):
        frame = _frame(pkg, gold, i, nl, thr, w, h)
        gk, gd = canon(gold[f"c{i}_kp"], gold[f"c{i}_desc"])
        with pkg.Context(max_width=w, max_height=h, max_batch=3, nfeatures=nf, nlevels=nl, scale_factor=sf, fast_threshold=thr,
                         edge_threshold=edge) as ctx:
            kp, desc = ctx.detect(frame)
            _same(kp, desc, gk.astype(oracle.KP_DTYPE), gd, f"CUDA single frame, case {i}")
            # the batch structure (per-level FAST launches, no kernel graph) on the same frame
            kps, descs, cnt = ctx.detect_batch(np.stack([frame, frame, frame]))
            for b in range(3):
                _same(kps[b, :cnt[b]], descs[b, :cnt[b]], gk.astype(oracle.KP_DTYPE), gd, f"CUDA batch of 3, frame {b}, case {i}")
