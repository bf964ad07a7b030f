"""GPU parity tests of the second detector (DistributedOrbOpenCvDetector), through the C ABI
(mslam_b200_dist_*) against oracle/dist_orb_oracle.c.  Bar: bit-exact at every stage -- pyramid levels, blurred levels,
per-level candidate / kept counts and tree passes -- and bit-exact keypoints (x, y, size, angle, response, octave as raw
bytes) and descriptors IN THE REFERENCE'S ORDER (index i is the reference's Keypoint::id).  Nothing reads /root/reference."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _check_frame(pkg, oracle, ctx, bgr, frame=0, kp=None, desc=None, what="", params=None):
    H, W, _ = bgr.shape
    okp, odesc, dumps = oracle.dist_detect_gray(oracle.gray(bgr), params, dumps=True)
    nl = params.nlevels if params else 8
    w, h, s = ctx.level_geometry(W, H)
    ow, oh, osc = oracle.dist_level_geometry(W, H, nl, params.scale_factor if params else 1.2)
    assert (w == ow).all() and (h == oh).all() and s.tobytes() == osc.tobytes()
    for l in range(nl):
        got = ctx.debug_image(pkg.DIST_PYRAMID, frame, l, W, H)
        assert (got == dumps["pyr"][l]).all(), f"{what} pyramid level {l}: {(got != dumps['pyr'][l]).sum()} px differ"
        assert ctx.debug_count(pkg.DIST_N_FAST, frame, l) == dumps["n_fast"][l], f"{what} n_fast level {l}"
        assert ctx.debug_count(pkg.DIST_N_TREE, frame, l) == dumps["n_tree"][l], f"{what} n_tree level {l}"
        if dumps["n_fast"][l] > 0:
            assert ctx.debug_count(pkg.DIST_PASSES, frame, l) == dumps["passes"][l], f"{what} passes level {l}"
        if dumps["n_tree"][l] > 0:  # the reference blurs only levels that kept keypoints (:789-799)
            got = ctx.debug_image(pkg.DIST_BLURRED, frame, l, W, H)
            assert (got == dumps["blurred"][l]).all(), f"{what} blurred level {l}"
    if kp is not None:
        assert len(kp) == len(okp), f"{what}: {len(kp)} keypoints vs oracle {len(okp)}"
        if kp.tobytes() != okp.tobytes():
            bad = [i for i in range(len(kp)) if kp[i].tobytes() != okp[i].tobytes()]
            raise AssertionError(f"{what}: {len(bad)} keypoints differ, first {bad[0]}: {kp[bad[0]]} vs {okp[bad[0]]}")
        nbad = int((desc != odesc).any(axis=1).sum())
        assert nbad == 0, f"{what}: {nbad}/{len(desc)} descriptors differ"
    return okp, odesc


def test_bundled_frames_bit_exact_in_reference_order(pkg, oracle, bundled):
    with pkg.DistContext(max_width=640, max_height=480, max_batch=2) as ctx:
        for name in ("bgr0", "bgr1"):
            kp, desc = ctx.detect(bundled[name])
            assert len(kp) > 1000
            _check_frame(pkg, oracle, ctx, bundled[name], 0, kp, desc, name)


@pytest.mark.parametrize("W,H", [(640, 480), (1280, 720), (1920, 1080), (311, 237), (480, 640), (333, 500), (160, 140)])
def test_synthetic_sizes_bit_exact(pkg, oracle, W, H):
    frames, _ = pkg.synth.frame_batch(W, H, 1, 0, seed=W + H)
    with pkg.DistContext(max_width=W, max_height=H, max_batch=1) as ctx:
        kp, desc = ctx.detect(frames[0])
        _check_frame(pkg, oracle, ctx, frames[0], 0, kp, desc, f"{W}x{H}")


@pytest.mark.parametrize("nlevels,sf,ini,mn,min_size,W,H", [(5, 1.5, 30, 10, 400, 640, 480), (3, 2.0, 12, 12, 5000, 800, 300),
                                                            (8, 1.2, 20, 7, 150, 400, 400), (11, 1.1, 40, 5, 1000, 512, 384),
                                                            (3, 2.5, 20, 7, 1000, 900, 700)])
def test_non_default_parameters(pkg, oracle, nlevels, sf, ini, mn, min_size, W, H):
    """orb_params / min_size other than the hard-coded ones, including an exact 2x pyramid, a 2.5x one (the resize kernel's
    per-pixel fallback: four output pixels span more than the quad kernel's 8-byte window) and several initial tree nodes"""
    frames, _ = pkg.synth.frame_batch(W, H, 1, 0, seed=nlevels * 100 + ini)
    prm = oracle.dist_default_params(nlevels=nlevels, ini_fast_thr=ini, min_fast_thr=mn, min_size=min_size, scale_factor=sf)
    with pkg.DistContext(max_width=W, max_height=H, nlevels=nlevels, scale_factor=sf, ini_fast_threshold=ini,
                         min_fast_threshold=mn, min_size=min_size) as ctx:
        kp, desc = ctx.detect(frames[0])
        assert len(kp) > 0
        _check_frame(pkg, oracle, ctx, frames[0], 0, kp, desc, f"{nlevels} levels x{sf}", params=prm)


def test_noise_frame_many_candidates(pkg, oracle):
    rng = np.random.default_rng(11)
    bgr = rng.integers(0, 256, (480, 640, 3), dtype=np.uint8)
    with pkg.DistContext(max_width=640, max_height=480) as ctx:
        kp, desc = ctx.detect(bgr)
        _check_frame(pkg, oracle, ctx, bgr, 0, kp, desc, "noise")


def test_low_contrast_uses_the_threshold_fallback(pkg, oracle):
    """contrast between 7 and 20: no corner at ini_fast_thr_, so every cell is redone at min_fast_thr_ (:919-923)"""
    rng = np.random.default_rng(5)
    g = (120 + rng.integers(0, 14, (480, 640))).astype(np.uint8)
    bgr = np.repeat(g[:, :, None], 3, axis=2)
    with pkg.DistContext(max_width=640, max_height=480) as ctx:
        kp, desc = ctx.detect(bgr)
        okp, _ = _check_frame(pkg, oracle, ctx, bgr, 0, kp, desc, "low contrast")
        assert len(okp) > 0 and okp["response"].max() < 20


def test_flat_frame_has_no_keypoints(pkg, oracle):
    bgr = np.full((240, 320, 3), 77, np.uint8)
    with pkg.DistContext(max_width=320, max_height=240) as ctx:
        kp, desc = ctx.detect(bgr)
        assert len(kp) == 0 and len(desc) == 0


def test_batch_of_frames_equals_single_frames(pkg, oracle):
    frames, _ = pkg.synth.frame_batch(640, 480, 5, 1, seed=77)
    with pkg.DistContext(max_width=640, max_height=480, max_batch=5) as ctx:
        kp, desc, counts = ctx.detect_batch(frames)
        for i in range(5):
            _check_frame(pkg, oracle, ctx, frames[i], i, kp[i, :counts[i]], desc[i, :counts[i]], f"frame {i}")


def test_device_entry_point(pkg, oracle):
    import torch
    frames, _ = pkg.synth.frame_batch(640, 480, 3, 0, seed=9)
    with pkg.DistContext(max_width=640, max_height=480, max_batch=3) as ctx:
        cap = ctx.kp_cap
        d_bgr = torch.from_numpy(frames).cuda()
        d_kp = torch.zeros((3, cap, 6), dtype=torch.float32, device="cuda")
        d_desc = torch.zeros((3, cap, 32), dtype=torch.uint8, device="cuda")
        d_n = torch.zeros(3, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        ctx.detect_batch_dev(d_bgr, 3, 640, 480, d_kp, d_desc, cap, d_n)
        ctx.synchronize()
        n = d_n.cpu().numpy()
        kp = d_kp.cpu().numpy().view(np.uint8).reshape(3, cap, 24).copy().view(pkg.KP_DTYPE).reshape(3, cap)
        desc = d_desc.cpu().numpy()
        for i in range(3):
            okp, odesc = oracle.dist_detect_rgb(frames[i])
            assert n[i] == len(okp)
            assert kp[i, :n[i]].tobytes() == okp.tobytes() and (desc[i, :n[i]] == odesc).all()


def test_frontend_class_sets_ids(pkg, oracle, bundled):
    det = pkg.DistributedOrbB200Detector(max_width=640, max_height=480)
    out = det.detect(pkg.RgbFrame(bundled["bgr0"], 640, 480))
    okp, odesc = oracle.dist_detect_rgb(bundled["bgr0"])
    assert (out["id"] == np.arange(len(okp))).all()
    assert (out["coordinates"][:, 0] == okp["x"].astype(np.float64)).all()
    assert (out["coordinates"][:, 1] == okp["y"].astype(np.float64)).all()
    assert (out["descriptor"] == odesc).all()
    det.ctx.close()


def test_capacity_and_invalid_sizes(pkg):
    frames, _ = pkg.synth.frame_batch(640, 480, 1, 0, seed=3)
    with pkg.DistContext(max_width=640, max_height=480) as ctx:
        with pytest.raises(pkg.MslamB200Error) as e:
            ctx.detect(frames[0], capacity=10)
        assert e.value.status == pkg.E_CAPACITY
        with pytest.raises(pkg.MslamB200Error) as e:
            ctx.detect(np.zeros((481, 640, 3), np.uint8))  # larger than the handle
        assert e.value.status == pkg.E_INVALID
        with pytest.raises(pkg.MslamB200Error) as e:
            ctx.detect(np.zeros((100, 100, 3), np.uint8))  # level 7 is 28 px: the reference's border arithmetic underflows
        assert e.value.status == pkg.E_INVALID


def test_pipelined_host_batches_equal_blocking_calls(pkg):
    B, W, H = 8, 320, 240
    batches, bases = [], None
    for t in range(4):
        fr, bases = pkg.synth.frame_batch(W, H, B, t, seed=33, bases=bases)
        batches.append(fr)
    with pkg.DistContext(max_width=W, max_height=H, max_batch=B) as ctx:
        cap = ctx.kp_cap
        ref = [ctx.detect_batch(f) for f in batches]
        hf = []
        for f in batches:
            h = pkg.pinned_empty(f.shape, np.uint8)
            h[...] = f
            hf.append(h)
        outs = [(pkg.pinned_empty((B, cap), pkg.KP_DTYPE), pkg.pinned_empty((B, cap, 32), np.uint8), pkg.pinned_empty((B,), np.int32))
                for _ in range(4)]
        ctx.detect_batch_begin(hf[0], cap, outs[0])
        ctx.detect_batch_begin(hf[1], cap, outs[1])
        with pytest.raises(pkg.MslamB200Error):
            ctx.detect_batch_begin(hf[2], cap, outs[2])
        for t in range(4):
            kp, desc, n = ctx.detect_batch_end()
            if t + 2 < 4:
                ctx.detect_batch_begin(hf[t + 2], cap, outs[t + 2])
            assert (n == ref[t][2]).all() and n.min() > 0
            for b in range(B):
                assert kp[b, :n[b]].tobytes() == ref[t][0][b, :n[b]].tobytes() and (desc[b, :n[b]] == ref[t][1][b, :n[b]]).all()
        with pytest.raises(pkg.MslamB200Error):
            ctx.detect_batch_end()


def test_single_frame_graph_replay_matches_oracle(pkg, oracle):
    """Single-frame calls replay a captured kernel graph from the third call of a geometry on (per frame slot): a stream of
    different frames, a geometry change in the middle and a return to the first geometry all stay bit-exact, and the
    launch counter advances by the same amount for a replayed frame as for a directly launched one."""
    W, H = 320, 240
    frames, bases = [], None
    for t in range(7):
        fr, bases = pkg.synth.frame_batch(W, H, 1, t, seed=91, bases=bases)
        frames.append(fr[0])
    other, _ = pkg.synth.frame_batch(400, 300, 1, 0, seed=92)
    with pkg.DistContext(max_width=400, max_height=300, max_batch=1) as ctx:
        per_frame = []
        for t, f in enumerate(frames):
            l0 = ctx.launch_count
            kp, desc = ctx.detect(f)
            per_frame.append(ctx.launch_count - l0)
            _check_frame(pkg, oracle, ctx, f, 0, kp, desc, f"frame {t}")
            if t == 4:
                kp, desc = ctx.detect(other[0])
                _check_frame(pkg, oracle, ctx, other[0], 0, kp, desc, "other geometry")
        assert len(set(per_frame)) == 1 and per_frame[0] >= 10
