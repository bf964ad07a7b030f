"""GPU parity tests (run with -m gpu on the B200 box): the CUDA path, called through the C ABI of
include/mslam_b200.h, against the CPU oracle (oracle/orb_oracle.c) and the committed cv2 golden vectors.

Bar (BASELINE.json north_star): descriptors, Hamming distances and match indices bit-exact; keypoints are
compared bit-exact too (x, y, size, angle, response, octave as f32/i32 bytes) -- the stated fallback
tolerances (0.5 px, response rel 1e-6, angle 1e-3 deg, overlap >= 99 %) are not needed.
Nothing here reads /root/reference.
"""
import numpy as np
import pytest

from conftest import canon

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx(pkg):
    with pkg.Context(max_width=640, max_height=480, max_batch=4, nfeatures=1000) as c:
        yield c


def _assert_same_keypoints(kp, desc, okp, odesc, what=""):
    assert len(kp) == len(okp), f"{what}: {len(kp)} keypoints vs oracle {len(okp)}"
    if kp.tobytes() != okp.tobytes():
        bad = [i for i in range(len(kp)) if kp[i].tobytes() != okp[i].tobytes()]
        raise AssertionError(f"{what}: {len(bad)} keypoints differ, first {bad[0]}: {kp[bad[0]]} vs {okp[bad[0]]}")
    nbad = int((desc != odesc).any(axis=1).sum())
    assert nbad == 0, f"{what}: {nbad}/{len(desc)} descriptors differ"


def _bgr_with_gray(oracle, g):
    """a 3-channel image whose reference gray conversion (frame.cpp:18-19) is exactly `g`: (v, v, v) truncates to v - 1
    for some v (.299f + .587f + .114f products round down), so pick a fixed-up triple per gray level"""
    lut = np.zeros((256, 3), np.uint8)
    for v in range(256):
        for cand in ((v, v, v), (min(v + 1, 255), v, v), (v, min(v + 1, 255), v), (min(v + 1, 255), min(v + 1, 255), v)):
            if oracle.gray(np.array([[cand]], np.uint8))[0, 0] == v:
                lut[v] = cand
                break
        else:
            raise AssertionError(v)
    bgr = lut[g]
    assert (oracle.gray(bgr) == g).all()
    return bgr


# ---- stage by stage ----------------------------------------------------------------------------------------------------
def test_gray_pyramid_and_blur_levels_bit_exact(pkg, ctx, oracle, bundled):
    bgr = bundled["bgr0"]
    ctx.detect(bgr)
    _, _, dumps = oracle.orb_detect_gray(oracle.gray(bgr), oracle.default_params(1000), dumps=True)
    w, h, s, nl = ctx.level_geometry(640, 480)
    for l in range(8):
        got = ctx.debug_read(pkg.STAGE_PYRAMID, 0, l, int(w[l]), int(h[l]))
        assert got.shape == dumps["pyr"][l].shape
        assert (got == dumps["pyr"][l]).all(), f"pyramid level {l}: {(got != dumps['pyr'][l]).sum()} px differ"
        got = ctx.debug_read(pkg.STAGE_BLURRED, 0, l, int(w[l]), int(h[l]))
        assert (got == dumps["blurred"][l]).all(), f"blurred level {l}: {(got != dumps['blurred'][l]).sum()} px differ"
        assert ctx.debug_read(pkg.STAGE_N_FAST, 0, l) == dumps["n_fast"][l]
        assert ctx.debug_read(pkg.STAGE_N_CUT1, 0, l) == dumps["n_cut1"][l]


def test_level_geometry_matches_oracle(ctx, oracle):
    for (W, H) in ((640, 480), (311, 237), (1280, 720), (1920, 1080), (97, 65)):
        with_ctx = ctx.level_geometry(W, H)
        w, h, s = oracle.level_geometry(W, H)
        assert (with_ctx[0] == w).all() and (with_ctx[1] == h).all() and with_ctx[2].tobytes() == s.tobytes()
    assert (ctx.level_geometry(640, 480)[3] == oracle.features_per_level(1000)).all()


# ---- detect --------------------------------------------------------------------------------------------------------------
def test_bundled_frames_bit_exact_vs_cv2_golden(ctx, oracle, bundled):
    for i in (0, 1):
        kp, desc = ctx.detect(bundled[f"bgr{i}"])
        gk, gd = canon(bundled[f"kp{i}"], bundled[f"desc{i}"])
        _assert_same_keypoints(kp, desc, gk.astype(oracle.KP_DTYPE), gd, f"bundled frame {i}")
        assert np.bincount(kp["octave"]).tolist() == [217, 181, 151, 126, 105, 87, 73, 60]


def test_odd_size_noise_frame_vs_cv2_golden(pkg, oracle, synthetic):
    g = synthetic["noise_gray"]  # 237 x 311, level sizes hit the round-half-even case
    bgr = _bgr_with_gray(oracle, g)
    with pkg.Context(max_width=311, max_height=237, nfeatures=300) as c:
        kp, desc = c.detect(bgr)
    gk, gd = canon(synthetic["noise_kp"], synthetic["noise_desc"])
    _assert_same_keypoints(kp, desc, gk.astype(oracle.KP_DTYPE), gd, "noise frame")


@pytest.mark.parametrize("W,H,n", [(640, 480, 1000), (1280, 720, 2000), (1920, 1080, 4000), (322, 250, 500)])
def test_synthetic_frames_vs_oracle(pkg, oracle, W, H, n):
    frames, bases = pkg.synth.frame_batch(W, H, 2, 0, seed=77 + W)
    with pkg.Context(max_width=W, max_height=H, max_batch=2, nfeatures=n) as c:
        kp, desc, cnt = c.detect_batch(frames)
        for b in range(2):
            okp, odesc = oracle.orb_detect_rgb(frames[b], oracle.default_params(n))
            _assert_same_keypoints(kp[b, :cnt[b]], desc[b, :cnt[b]], okp, odesc, f"{W}x{H} frame {b}")
            if W >= 640:  # the generator must make retainBest bite on every level at the BASELINE.json sizes
                assert cnt[b] >= n
                for l in range(8):
                    assert c.debug_read(pkg.STAGE_N_FAST, b, l) > 2 * c.level_geometry(W, H)[3][l]


@pytest.mark.parametrize("W,H,n", [(1280, 720, 6000), (1920, 1080, 10000)])
def test_feature_counts_beyond_the_selection_kernels_shared_lists(pkg, oracle, W, H, n):
    """k_select keeps the survivors of both retainBest cuts in shared memory up to 2048 entries per level and falls back to
    its global lists beyond: 6000 features put level 0's first cut (2 n_0 + ties) past that, 10000 the final list as well"""
    frames, _ = pkg.synth.frame_batch(W, H, 1, 0, seed=5 + W)
    with pkg.Context(max_width=W, max_height=H, max_batch=1, nfeatures=n) as c:
        n0 = int(c.level_geometry(W, H)[3][0])
        assert 2 * n0 > 2048 and (n < 10000 or n0 > 2048)
        kp, desc = c.detect(frames[0])
        assert c.debug_read(pkg.STAGE_N_CUT1, 0, 0) > 2048
        okp, odesc = oracle.orb_detect_rgb(frames[0], oracle.default_params(n))
        _assert_same_keypoints(kp, desc, okp, odesc, f"{W}x{H} / {n} features")


def test_batch_equals_single(pkg, ctx):
    frames, _ = pkg.synth.frame_batch(640, 480, 4, 1, seed=5)
    kp, desc, cnt = ctx.detect_batch(frames)
    for b in range(4):
        k1, d1 = ctx.detect(frames[b])
        assert cnt[b] == len(k1) and kp[b, :cnt[b]].tobytes() == k1.tobytes() and (desc[b, :cnt[b]] == d1).all()


def test_flat_and_tiny_frames(pkg, oracle):
    with pkg.Context(max_width=640, max_height=480, nfeatures=1000) as c:
        kp, desc = c.detect(np.full((480, 640, 3), 128, np.uint8))
        assert len(kp) == 0
        small = np.random.default_rng(0).integers(0, 256, (70, 90, 3), dtype=np.uint8)  # most levels < 2*edge
        kp, desc = c.detect(small)
        okp, odesc = oracle.orb_detect_rgb(small, oracle.default_params(1000))
        _assert_same_keypoints(kp, desc, okp, odesc, "tiny frame")


def test_row_stride_and_capacity(pkg, ctx, bundled):
    bgr = bundled["bgr0"]
    wide = np.zeros((480, 700, 3), np.uint8)
    wide[:, :640] = bgr
    kp0, d0 = ctx.detect(bgr)
    kp, desc, cnt = ctx.detect_batch(wide[None, :, :640])  # row stride 2100 bytes
    assert cnt[0] == len(kp0) and kp[0, :cnt[0]].tobytes() == kp0.tobytes() and (desc[0, :cnt[0]] == d0).all()
    kp, desc, cnt = ctx.detect_batch(bgr[None], capacity=100, allow_capacity_error=True)
    assert ctx.last_status == pkg.E_CAPACITY and cnt[0] == 1000
    assert kp[0].tobytes() == kp0[:100].tobytes()


def test_describe_given_cv2_keypoints(ctx, oracle, bundled):
    # "descriptors bit-exact given identical keypoints": cv2's own keypoints (native order) in, descriptors out
    kps = bundled["kp0"].astype(oracle.KP_DTYPE)
    out, desc = ctx.describe(bundled["bgr0"], kps)
    assert (desc == bundled["desc0"]).all()
    assert out["angle"].tobytes() == kps["angle"].tobytes()


def test_detect_dev_entry_point(pkg, ctx, bundled):
    import torch
    bgr = torch.from_numpy(np.stack([bundled["bgr0"], bundled["bgr1"]])).cuda()
    cap = ctx.kp_cap
    d_kp = torch.zeros((2, cap, 6), dtype=torch.float32, device="cuda")
    d_desc = torch.zeros((2, cap, 32), dtype=torch.uint8, device="cuda")
    d_cnt = torch.zeros(2, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    ctx.detect_batch_dev(bgr, 2, 640, 480, d_kp, d_desc, cap, d_cnt)
    ctx.synchronize()
    for i in (0, 1):
        k1, d1 = ctx.detect(bundled[f"bgr{i}"])
        n = int(d_cnt[i])
        assert n == len(k1)
        assert d_kp[i, :n].cpu().numpy().tobytes() == k1.tobytes()
        assert (d_desc[i, :n].cpu().numpy() == d1).all()


# ---- match ---------------------------------------------------------------------------------------------------------------
def test_bundled_matches_vs_cv2_golden(ctx, bundled):
    d0, d1 = bundled["desc0"], bundled["desc1"]
    fi, ti, di = ctx.match_knn_ratio(d0, d1)
    g = bundled["match_0_to_1"]
    assert len(fi) == 693
    assert (np.stack([fi.astype(np.int64), ti.astype(np.int64), di], 1) == g).all()
    qi, tj, dj = ctx.match_crosscheck(d1, d0)
    assert len(qi) == 765 and (np.stack([qi, tj, dj], 1) == bundled["cross_q1_t0"]).all()
    i0, e0, i1, e1 = ctx.knn2(d1, d0)
    assert (np.stack([i0, e0, i1, e1], 1) == bundled["knn2_q1_t0"]).all()


def test_strided_orb_keypoint_records(pkg, ctx, bundled):
    # orb_feature.cpp:88-91: descriptors viewed in place at stride sizeof(OrbKeypoint) = 64, offset 32
    rec0 = np.zeros(1000, pkg.frontend.ORB_KEYPOINT_DTYPE)
    rec1 = np.zeros(1000, pkg.frontend.ORB_KEYPOINT_DTYPE)
    rec0["descriptor"], rec1["descriptor"] = bundled["desc0"], bundled["desc1"]
    fi, ti, di = ctx.match_knn_ratio(rec0["descriptor"], rec1["descriptor"])
    assert (np.stack([fi.astype(np.int64), ti.astype(np.int64), di], 1) == bundled["match_0_to_1"]).all()


def test_tie_heavy_sets_vs_cv2_golden(ctx, synthetic):
    a, b = synthetic["tie_a"], synthetic["tie_b"]
    i0, e0, i1, e1 = ctx.knn2(a, b)
    assert (np.stack([i0, e0, i1, e1], 1) == synthetic["tie_knn2"]).all()
    qi, tj, dj = ctx.match_crosscheck(a, b)
    assert (np.stack([qi, tj, dj], 1) == synthetic["tie_cross"]).all()
    fi, ti, di = ctx.match_knn_ratio(b, a)
    assert (np.stack([fi.astype(np.int64), ti.astype(np.int64), di], 1) == synthetic["tie_ratio"]).all()


def test_matcher_edge_cases(ctx, oracle):
    rng = np.random.default_rng(0)
    d = rng.integers(0, 256, (5, 32), dtype=np.uint8)
    empty = np.zeros((0, 32), np.uint8)
    assert len(ctx.match_knn_ratio(empty, d)[0]) == 0
    assert len(ctx.match_knn_ratio(d, empty)[0]) == 0
    assert len(ctx.match_knn_ratio(d[:1], d)[0]) == 0
    assert len(ctx.match_crosscheck(empty, d)[0]) == 0
    fi, ti, di = ctx.match_knn_ratio(d, d)
    assert (fi == ti).all() and (di == 0).all() and len(fi) == 5
    i0, e0, i1, e1 = ctx.knn2(d, d[:1])  # train smaller than k
    assert (i0 == 0).all() and (i1 == -1).all() and (e1 == -1).all()
    for n_from, n_to in ((2, 1), (3, 1314), (1314, 3), (513, 129), (1000, 1001)):
        f = rng.integers(0, 256, (n_from, 32), dtype=np.uint8)
        t = rng.integers(0, 256, (n_to, 32), dtype=np.uint8)
        t[: min(n_from, n_to) // 2] = f[: min(n_from, n_to) // 2]  # true matches
        got = ctx.match_knn_ratio(f, t)
        ref = oracle.match_ratio(f, t)
        assert all((np.asarray(g).astype(np.int64) == r).all() for g, r in zip(got, ref)), (n_from, n_to)
        got = ctx.match_crosscheck(t, f)
        ref = oracle.match_crosscheck(t, f)
        assert all((g == r).all() for g, r in zip(got, ref)), (n_from, n_to)


def test_match_batch_ragged(ctx, oracle):
    rng = np.random.default_rng(3)
    P, cap = 4, ctx.desc_cap
    frm = rng.integers(0, 256, (P, cap, 32), dtype=np.uint8)
    to = rng.integers(0, 256, (P, cap, 32), dtype=np.uint8)
    nf = np.array([1000, 0, 777, cap], np.int32)
    nt = np.array([1000, 500, 1, cap], np.int32)
    for p in range(P):
        m = min(nf[p], nt[p]) // 2
        bits = np.unpackbits(frm[p, :m], axis=-1)
        to[p, :m] = np.packbits(bits ^ (rng.random(bits.shape) < 0.04), axis=-1)
    oa, ob, od, no = ctx.match_batch(0, frm, nf, to, nt)
    for p in range(P):
        fi, ti, di = oracle.match_ratio(frm[p, :nf[p]], to[p, :nt[p]])
        assert no[p] == len(fi)
        assert (oa[p, :no[p]] == fi).all() and (ob[p, :no[p]] == ti).all() and (od[p, :no[p]] == di).all()
    oa, ob, od, no = ctx.match_batch(1, frm, nf, to, nt)
    for p in range(P):
        qi, tj, dj = oracle.match_crosscheck(to[p, :nt[p]], frm[p, :nf[p]])
        assert no[p] == len(qi)
        assert (ob[p, :no[p]] == qi).all() and (oa[p, :no[p]] == tj).all() and (od[p, :no[p]] == dj).all()


def test_full_size_match_properties(pkg):
    # 4000 x 4000 (configs[3]): properties that need no oracle
    rng = np.random.default_rng(11)
    d = rng.integers(0, 256, (4000, 32), dtype=np.uint8)
    with pkg.Context(nfeatures=4000, max_descriptors=4096) as c:
        fi, ti, di = c.match_knn_ratio(d, d)
        assert len(fi) == 4000 and (fi == ti).all() and (di == 0).all()
        perm = rng.permutation(4000)
        fi, ti, di = c.match_knn_ratio(d, d[perm])  # to[j] = from[perm[j]]
        assert (fi.astype(np.int64) == perm).all() and (ti == np.arange(4000)).all()
        qi, tj, dj = c.match_crosscheck(d[perm], d)
        assert (tj == perm).all() and (dj == 0).all()
        i0, e0, i1, e1 = c.knn2(d, d)
        assert (i0 == np.arange(4000)).all() and (e0 == 0).all() and (e1 >= e0).all() and (i1 != i0).all()


# ---- keyframe database -------------------------------------------------------------------------------------------------
def test_db_query_vs_oracle(pkg, ctx, oracle):
    rng = np.random.default_rng(5)
    q = rng.integers(0, 256, (5, 300, 32), dtype=np.uint8)
    db = pkg.synth.descriptor_db(40, 300, seed=9)
    planted = pkg.synth.plant_copies(db, q, frac=0.2, seed=4)
    qc = np.array([300, 300, 257, 2, 1], np.int32)
    kc = np.full(40, 300, np.int32)
    kc[3], kc[7] = 0, 33
    s, ids, allsc = ctx.db_query(q, qc, db, kc, k=5, id_base=1000, want_scores=True)
    ref = oracle.db_scores(q, qc, db, kc, nthreads=4)
    assert (allsc == ref).all()
    rs, rid = oracle.db_topk(ref, 5, id_base=1000)
    assert (s == rs).all() and (ids == rid).all()
    for kf, qi in planted.items():
        if qc[qi] == 300 and kc[kf] == 300:
            assert allsc[qi, kf] > 200  # a planted copy is found
    # k larger than the shard pads with (-1, -1)
    s, ids = ctx.db_query(q, qc, db[:3], kc[:3], k=5)
    assert (s[:, 3:] == -1).all() and (ids[:, 3:] == -1).all()


def test_db_sharded_merge_equals_single(pkg, ctx, oracle):
    """the N-GPU path emulated on one GPU: per-shard top-k lists merged == top-k of the whole DB"""
    import torch
    rng = np.random.default_rng(8)
    Q, K, n, k, shards = 6, 48, 128, 4, 4
    q = rng.integers(0, 256, (Q, n, 32), dtype=np.uint8)
    db = pkg.synth.descriptor_db(K, n, seed=2)
    pkg.synth.plant_copies(db, q, frac=0.25, seed=6)
    qc = np.full(Q, n, np.int32)
    kc = np.full(K, n, np.int32)
    whole_s, whole_i = ctx.db_query(q, qc, db, kc, k=k)
    per = K // shards
    lists_s = np.zeros((shards, Q, k), np.int32)
    lists_i = np.zeros((shards, Q, k), np.int32)
    for r in range(shards):
        lists_s[r], lists_i[r] = ctx.db_query(q, qc, db[r * per:(r + 1) * per], kc[r * per:(r + 1) * per], k=k,
                                              id_base=r * per)
    ds, di = torch.from_numpy(lists_s).cuda(), torch.from_numpy(lists_i).cuda()
    os_ = torch.zeros((Q, k), dtype=torch.int32, device="cuda")
    oi = torch.zeros((Q, k), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    ctx.db_merge_topk_dev(ds, di, shards, Q, k, os_, oi)
    ctx.synchronize()
    assert (os_.cpu().numpy() == whole_s).all() and (oi.cpu().numpy() == whole_i).all()


def test_sharded_keyframe_db_host_class(pkg, ctx, oracle):
    """ShardedKeyframeDb at world 1 (device tensors in/out) == oracle; the N > 1 exchange is covered by tests/test_sharded_db.py"""
    import torch
    from modular_slam_b200.keyframe_db import ShardedKeyframeDb
    rng = np.random.default_rng(3)
    Q, K, n, k = 4, 30, 200, 3
    q = rng.integers(0, 256, (Q, n, 32), dtype=np.uint8)
    db = pkg.synth.descriptor_db(K, n, seed=12)
    pkg.synth.plant_copies(db, q, frac=0.2, seed=1)
    qc, kc = np.full(Q, n, np.int32), np.full(K, n, np.int32)
    kdb = ShardedKeyframeDb(ctx)
    kdb.load_shard(torch.from_numpy(db).cuda(), torch.from_numpy(kc).cuda(), id_base=500)
    torch.cuda.synchronize()
    s, ids = kdb.query(torch.from_numpy(q).cuda(), torch.from_numpy(qc).cuda(), k)
    ctx.synchronize()
    rs, rid = oracle.db_topk(oracle.db_scores(q, qc, db, kc, nthreads=4), k, id_base=500)
    assert (s.cpu().numpy() == rs).all() and (ids.cpu().numpy() == rid).all()


# ---- the reference-facing mirror -------------------------------------------------------------------------------------
def test_frontend_mirror_reads_like_the_reference(pkg, bundled):
    det = pkg.OrbB200Detector(max_width=640, max_height=480)
    mat = pkg.OrbB200Matcher(ctx=det.ctx)
    f0 = det.detect(pkg.RgbFrame(bundled["bgr0"], 640, 480))
    f1 = det.detect(pkg.RgbFrame(bundled["bgr1"], 640, 480))
    assert f0.dtype.itemsize == 64 and len(f0) == 1000
    m = mat.match(f0, f1)
    # canonical keypoint order differs from cv2's, so compare through coordinates
    g0, g1 = bundled["kp0"], bundled["kp1"]
    gold = {(g0["x"][a], g0["y"][a], g1["x"][b], g1["y"][b]) for a, b, _ in bundled["match_0_to_1"]}
    got = {(np.float32(f0["coordinates"][a][0]), np.float32(f0["coordinates"][a][1]),
            np.float32(f1["coordinates"][b][0]), np.float32(f1["coordinates"][b][1]))
           for a, b in zip(m["fromIndex"], m["toIndex"])}
    # duplicated keypoint coordinates across octaves can alias; identical sets is the expectation
    assert len(m) == 693 and got == gold
    det.ctx.close()


def test_launch_counter_counts(ctx, bundled):
    n0 = ctx.launch_count
    ctx.detect(bundled["bgr0"])
    assert ctx.launch_count - n0 >= 5


def test_pipelined_host_batches_equal_blocking_calls(pkg):
    """mslam_b200_detect_batch_begin / _end with two batches in flight + match_batch_begin / _end == the blocking calls"""
    B, W, H = 8, 320, 240
    batches, bases = [], None
    for t in range(5):
        fr, bases = pkg.synth.frame_batch(W, H, B, t, seed=21, bases=bases)
        batches.append(fr)
    with pkg.Context(max_width=W, max_height=H, max_batch=B, nfeatures=300) as ctx:
        cap = ctx.kp_cap
        ref = [ctx.detect_batch(f, cap) for f in batches]
        ref_m = [ctx.match_batch(0, ref[t][1], ref[t][2], ref[t - 1][1], ref[t - 1][2]) for t in range(1, 5)]
        hf = []
        for f in batches:
            h = pkg.pinned_empty(f.shape, np.uint8)
            h[...] = f
            hf.append(h)
        outs = [(pkg.pinned_empty((B, cap), pkg.KP_DTYPE), pkg.pinned_empty((B, cap, 32), np.uint8), pkg.pinned_empty((B,), np.int32))
                for _ in range(5)]
        hm = [pkg.pinned_empty((B, cap), np.int32) for _ in range(3)] + [pkg.pinned_empty((B,), np.int32)]
        ctx.detect_batch_begin(hf[0], cap, outs[0])
        ctx.detect_batch_begin(hf[1], cap, outs[1])
        with pytest.raises(pkg.MslamB200Error):  # a third batch does not fit
            ctx.detect_batch_begin(hf[2], cap, outs[2])
        with pytest.raises(pkg.MslamB200Error):  # and the blocking call refuses to jump the queue
            ctx.detect_batch(batches[2], cap)
        for t in range(5):
            kp, desc, n = ctx.detect_batch_end()
            if t + 2 < 5:
                ctx.detect_batch_begin(hf[t + 2], cap, outs[t + 2])
            assert (n == ref[t][2]).all()
            for b in range(B):
                assert kp[b, :n[b]].tobytes() == ref[t][0][b, :n[b]].tobytes() and (desc[b, :n[b]] == ref[t][1][b, :n[b]]).all()
            if t >= 1:
                ctx.match_batch_begin(0, outs[t][1], outs[t][2], outs[t - 1][1], outs[t - 1][2], tuple(hm))
                oa, ob, od, no = ctx.match_batch_end()
                ra, rb, rd, rn = ref_m[t - 1]
                assert (no == rn).all()
                for b in range(B):
                    assert (oa[b, :no[b]] == ra[b, :rn[b]]).all() and (ob[b, :no[b]] == rb[b, :rn[b]]).all()
        with pytest.raises(pkg.MslamB200Error):
            ctx.detect_batch_end()  # nothing in flight


def test_single_frame_graph_replay_matches_oracle(pkg, oracle, bundled):
    """mslam_b200_detect replays a captured kernel graph from the third call of a geometry on (per frame slot): a stream
    of different frames, a geometry change in the middle, a describe-only call and a batch in between all stay
    bit-exact, and a replayed frame advances the launch counter like a directly launched one."""
    W, H = 320, 240
    frames, bases = [], None
    for t in range(7):
        fr, bases = pkg.synth.frame_batch(W, H, 1, t, seed=19, bases=bases)
        frames.append(fr[0])
    prm = oracle.default_params(500)
    with pkg.Context(max_width=640, max_height=480, max_batch=2, nfeatures=500) as c:
        per_frame = []
        for t, f in enumerate(frames):
            l0 = c.launch_count
            kp, desc = c.detect(f)
            per_frame.append(c.launch_count - l0)
            okp, odesc = oracle.orb_detect_rgb(f, prm)
            _assert_same_keypoints(kp, desc, okp, odesc, f"frame {t}")
            if t == 3:
                kp, desc = c.detect(bundled["bgr0"])
                okp, odesc = oracle.orb_detect_rgb(bundled["bgr0"], prm)
                _assert_same_keypoints(kp, desc, okp, odesc, "other geometry")
            if t == 5:
                kb, db, cnt = c.detect_batch(np.stack([frames[0], frames[1]]))
                okp, odesc = oracle.orb_detect_rgb(frames[1], prm)
                _assert_same_keypoints(kb[1, :cnt[1]], db[1, :cnt[1]], okp, odesc, "batch between replays")
        assert len(set(per_frame)) == 1 and per_frame[0] >= 10


@pytest.mark.parametrize("W,H,B", [(320, 240, 16), (640, 480, 24)])
def test_large_batches_take_the_fused_pyramid_tail_and_stay_bit_exact(pkg, oracle, W, H, B):
    """batches of >= 16 frames build their small pyramid levels in one cluster-per-frame launch (k_resize_tail): every level of
    every frame still equals the oracle's chained INTER_LINEAR_EXACT resize, and so do the keypoints and descriptors"""
    frames, _ = pkg.synth.frame_batch(W, H, B, 0, seed=W + B)
    with pkg.Context(max_width=W, max_height=H, max_batch=B, nfeatures=500) as c:
        kps, descs, counts = c.detect_batch(frames)
        w, h, _, _ = c.level_geometry(W, H)
        for b in (0, 7, B - 1):
            okp, od, dumps = oracle.orb_detect_gray(oracle.gray(frames[b]), oracle.default_params(500), dumps=True)
            for l in range(8):
                got = c.debug_read(pkg.STAGE_PYRAMID, b, l, int(w[l]), int(h[l]))
                assert (got == dumps["pyr"][l]).all(), f"frame {b} level {l}: {(got != dumps['pyr'][l]).sum()} px differ"
            n = int(counts[b])
            assert n == len(okp) and kps[b, :n].tobytes() == okp.tobytes() and (descs[b, :n] == od).all(), f"frame {b}"
