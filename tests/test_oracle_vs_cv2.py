"""The oracle AND the CUDA path against OpenCV itself on the inputs bench.py runs (tests/golden/bench_synth.npz, made by
tests/golden/make_golden_bench.py with cv2 4.13.0 -- the reference pins OpenCV 4.8.1, conanfile.py:28):
synth.frame_batch(seed 1234) frames t = 0, 1 of sequence 0 at 640x480 / 1000 keypoints, 1280x720 / 2000, 1920x1080 / 4000.

Bar: bit-exact -- keypoints as raw bytes (x, y, size, angle, response as f32, octave as i32) in canonical (octave, y, x)
order, descriptors, and the kNN / ratio (orb_feature.cpp:96-114) / cross-check results computed on cv2's own descriptor
order.  The CPU half runs with -m "not gpu"; the GPU half calls through the C ABI.  Nothing here reads /root/reference.
"""
import hashlib
import os

import numpy as np
import pytest

from conftest import GOLDEN, canon

SHAPES = ((640, 480, 1000), (1280, 720, 2000), (1920, 1080, 4000))


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLDEN, "bench_synth.npz"))


@pytest.fixture(scope="module")
def frames(pkg, gold):
    """the bench generator's frames, regenerated (seeded numpy) and checked against the recorded sha256"""
    out = {}
    for w, h, _ in SHAPES:
        bases = None
        for t in (0, 1):
            f, bases = pkg.synth.frame_batch(w, h, 1, t, seed=1234, bases=bases)
            assert hashlib.sha256(f[0].tobytes()).hexdigest() == str(gold[f"{w}x{h}_sha{t}"]), \
                "synth.frame_batch changed: regenerate tests/golden/bench_synth.npz (make_golden_bench.py)"
            out[(w, h, t)] = f[0]
    return out


def _same(kp, desc, gk, gd, what):
    assert len(kp) == len(gk), f"{what}: {len(kp)} keypoints, cv2 has {len(gk)}"
    if kp.tobytes() != gk.tobytes():
        bad = [i for i in range(len(kp)) if kp[i].tobytes() != gk[i].tobytes()]
        raise AssertionError(f"{what}: {len(bad)} keypoints differ from cv2, first {bad[0]}: {kp[bad[0]]} vs {gk[bad[0]]}")
    nbad = int((desc != gd).any(axis=1).sum())
    assert nbad == 0, f"{what}: {nbad}/{len(desc)} descriptors differ from cv2"


@pytest.mark.parametrize("w,h,n", SHAPES)
def test_oracle_detect_equals_cv2_on_bench_frames(oracle, gold, frames, w, h, n):
    for t in (0, 1):
        kp, desc = oracle.orb_detect_rgb(frames[(w, h, t)], oracle.default_params(n))
        gk, gd = canon(gold[f"{w}x{h}_kp{t}"], gold[f"{w}x{h}_desc{t}"])
        _same(kp, desc, gk.astype(oracle.KP_DTYPE), gd, f"oracle {w}x{h} t={t}")
        assert np.bincount(kp["octave"]).tolist() == oracle.features_per_level(n).tolist()


@pytest.mark.parametrize("w,h,n", SHAPES)
def test_oracle_matchers_equal_cv2_on_bench_descriptors(oracle, gold, w, h, n):
    d0, d1 = gold[f"{w}x{h}_desc0"], gold[f"{w}x{h}_desc1"]  # cv2's own order, as the reference's matcher sees them
    fi, ti, di = oracle.match_ratio(d0, d1)
    assert (np.stack([fi, ti, di], 1) == gold[f"{w}x{h}_ratio_0_to_1"]).all()
    i0, e0, i1, e1 = oracle.knn2(d1, d0)
    assert (np.stack([i0, e0, i1, e1], 1) == gold[f"{w}x{h}_knn2_q1_t0"]).all()
    qi, tj, dj = oracle.match_crosscheck(d1, d0)
    assert (np.stack([qi, tj, dj], 1) == gold[f"{w}x{h}_cross_q1_t0"]).all()


def test_bench_goldens_are_in_the_tie_regime(gold):
    """the point of these fixtures: the synthetic stream has far more FAST corners than 2 n_l on every level and many
    equal integer scores, so retainBest()'s 'keep all ties' rule and the Harris cut are both exercised"""
    assert str(gold["cv2_version"]).startswith("4.")
    for w, h, n in SHAPES:
        kp = gold[f"{w}x{h}_kp0"]
        assert len(kp) == n and len(gold[f"{w}x{h}_ratio_0_to_1"]) > n // 2  # consecutive frames do match


# ---- the CUDA path, through the C ABI -------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("w,h,n", SHAPES)
def test_gpu_detect_equals_cv2_on_bench_frames(pkg, oracle, gold, frames, w, h, n):
    with pkg.Context(max_width=w, max_height=h, max_batch=2, nfeatures=n) as ctx:
        batch = np.stack([frames[(w, h, 0)], frames[(w, h, 1)]])
        kps, descs, counts = ctx.detect_batch(batch)
        for t in (0, 1):
            gk, gd = canon(gold[f"{w}x{h}_kp{t}"], gold[f"{w}x{h}_desc{t}"])
            c = int(counts[t])
            _same(kps[t, :c], descs[t, :c], gk.astype(oracle.KP_DTYPE), gd, f"gpu batch {w}x{h} t={t}")
            kp1, d1 = ctx.detect(frames[(w, h, t)])  # the single-frame (graph-replayed) call as well
            _same(kp1, d1, gk.astype(oracle.KP_DTYPE), gd, f"gpu single {w}x{h} t={t}")


@pytest.mark.gpu
@pytest.mark.parametrize("flags_name", ["FLAG_KNN_SCALAR", "FLAG_KNN_TENSOR"])
@pytest.mark.parametrize("w,h,n", SHAPES)
def test_gpu_matchers_equal_cv2_on_bench_descriptors(pkg, gold, w, h, n, flags_name):
    d0, d1 = gold[f"{w}x{h}_desc0"], gold[f"{w}x{h}_desc1"]
    with pkg.Context(max_width=64, max_height=64, max_batch=1, nfeatures=n, max_descriptors=n,
                     flags=getattr(pkg, flags_name)) as ctx:
        fi, ti, di = ctx.match_knn_ratio(d0, d1)
        g = gold[f"{w}x{h}_ratio_0_to_1"]
        assert (np.stack([fi.astype(np.int64), ti.astype(np.int64), di], 1) == g).all()
        i0, e0, i1, e1 = ctx.knn2(d1, d0)
        assert (np.stack([i0, e0, i1, e1], 1) == gold[f"{w}x{h}_knn2_q1_t0"]).all()
        qi, tj, dj = ctx.match_crosscheck(d1, d0)
        assert (np.stack([qi, tj, dj], 1) == gold[f"{w}x{h}_cross_q1_t0"]).all()
