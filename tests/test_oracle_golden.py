"""CPU: the oracle (oracle/orb_oracle.c) against the committed golden vectors produced by OpenCV itself
(tests/golden/make_golden.py, cv2 4.13.0 on the reference's bundled frames).  Bit-exact everywhere."""
import numpy as np

from conftest import canon


def test_gray_matches_frame_cpp_statement(oracle, bundled):
    bgr = bundled["bgr0"]
    f = bgr.astype(np.float32)
    ref = np.minimum(np.float32(255), (np.float32(0.299) * f[..., 0] + np.float32(0.587) * f[..., 1])
                     + np.float32(0.114) * f[..., 2]).astype(np.uint8)
    assert (oracle.gray(bgr) == ref).all()
    assert int(ref.sum()) == 32215673  # SURVEY 8(c) golden


def test_gray_saturates_and_truncates(oracle):
    px = np.array([[[255, 255, 255], [0, 0, 0], [1, 1, 1], [254, 255, 255], [10, 200, 30]]], np.uint8)
    assert oracle.gray(px).tolist() == [[255, 0, 1, 254, 123]]  # 255 -> min(255, 255.00002); 123.81 truncates


def test_bundled_keypoints_and_descriptors_bit_exact(oracle, bundled):
    for i in (0, 1):
        kp, desc = oracle.orb_detect_rgb(bundled[f"bgr{i}"])
        gk, gd = canon(bundled[f"kp{i}"], bundled[f"desc{i}"])
        assert len(kp) == 1000
        assert np.bincount(kp["octave"]).tolist() == [217, 181, 151, 126, 105, 87, 73, 60]
        assert kp.tobytes() == gk.astype(oracle.KP_DTYPE).tobytes()
        assert (desc == gd).all()


def test_bundled_matches(oracle, bundled):
    d0, d1 = bundled["desc0"], bundled["desc1"]
    fi, ti, di = oracle.match_ratio(d0, d1)
    g = bundled["match_0_to_1"]
    assert len(fi) == 693 and g.shape == (693, 3)
    assert (np.stack([fi, ti, di], 1) == g).all()
    assert g[:5].tolist() == [[154, 0, 9], [58, 1, 25], [127, 2, 14], [21, 4, 31], [52, 5, 16]]
    qi, tj, dj = oracle.match_crosscheck(d1, d0)
    assert len(qi) == 765 and (np.stack([qi, tj, dj], 1) == bundled["cross_q1_t0"]).all()
    i0, e0, i1, e1 = oracle.knn2(d1, d0)
    assert (np.stack([i0, e0, i1, e1], 1) == bundled["knn2_q1_t0"]).all()


def test_strided_descriptor_view_like_orb_keypoint(oracle, bundled):
    # orb_feature.cpp:88-91 views descriptors in place at stride sizeof(OrbKeypoint) = 64, offset 32
    rec0 = np.zeros((1000, 64), np.uint8)
    rec1 = np.zeros((1000, 64), np.uint8)
    rec0[:, 32:], rec1[:, 32:] = bundled["desc0"], bundled["desc1"]
    fi, ti, di = oracle.match_ratio(rec0[:, 32:], rec1[:, 32:])
    assert (np.stack([fi, ti, di], 1) == bundled["match_0_to_1"]).all()


def test_stage_primitives_on_noise(oracle, synthetic):
    g = synthetic["noise_gray"]
    score = oracle.fast_score_map(g, 20)
    got = oracle.fast_nms_border(score, edge=3)
    ref = synthetic["noise_fast"]
    got = got[np.lexsort((got[:, 0], got[:, 1]))]
    ref = ref[np.lexsort((ref[:, 0], ref[:, 1]))]
    assert got.shape == ref.shape and (got == ref).all()
    assert (oracle.resize_exact(g, 259, 198) == synthetic["noise_resize"]).all()
    assert (oracle.blur7(g) == synthetic["noise_blur"]).all()
    kp, desc = oracle.orb_detect_gray(g, oracle.default_params(300))
    gk, gd = canon(synthetic["noise_kp"], synthetic["noise_desc"])
    assert kp.tobytes() == gk.astype(oracle.KP_DTYPE).tobytes() and (desc == gd).all()


def test_tie_heavy_sets(oracle, synthetic):
    a, b = synthetic["tie_a"], synthetic["tie_b"]
    i0, e0, i1, e1 = oracle.knn2(a, b)
    assert (np.stack([i0, e0, i1, e1], 1) == synthetic["tie_knn2"]).all()
    qi, tj, dj = oracle.match_crosscheck(a, b)
    assert (np.stack([qi, tj, dj], 1) == synthetic["tie_cross"]).all()
    fi, ti, di = oracle.match_ratio(b, a)
    assert (np.stack([fi, ti, di], 1) == synthetic["tie_ratio"]).all()


def test_matcher_edge_cases(oracle):
    rng = np.random.default_rng(0)
    d = rng.integers(0, 256, (5, 32), dtype=np.uint8)
    empty = np.zeros((0, 32), np.uint8)
    assert len(oracle.match_ratio(empty, d)[0]) == 0
    assert len(oracle.match_ratio(d, empty)[0]) == 0
    assert len(oracle.match_ratio(d[:1], d)[0]) == 0  # from.size() < 2: reference reads match[1] out of range
    assert len(oracle.match_crosscheck(empty, d)[0]) == 0
    fi, ti, di = oracle.match_ratio(d, d)  # identical sets: d0 = 0 < 0.7 * d1 whenever d1 > 0
    assert (fi == ti).all() and (di == 0).all()


def test_integer_ratio_test_equals_reference_float_test():
    # orb_feature.cpp:101  m[0].distance < 0.7 * m[1].distance  (float distance, double product)
    for d0 in range(257):
        for d1 in range(257):
            assert (float(np.float32(d0)) < 0.7 * float(np.float32(d1))) == (10 * d0 < 7 * d1)


def test_level_geometry_and_budget(oracle):
    w, h, s = oracle.level_geometry(640, 480)
    assert list(zip(w.tolist(), h.tolist())) == [(640, 480), (533, 400), (444, 333), (370, 278), (309, 231), (257, 193),
                                                 (214, 161), (179, 134)]
    assert ["%08x" % v for v in s.view(np.uint32)] == ["3f800000", "3f99999a", "3fb851ec", "3fdd2f1c", "4004b5de",
                                                            "401f40a5", "403f1a60", "406552da"]
    assert oracle.features_per_level(1000).tolist() == [217, 181, 151, 126, 105, 87, 73, 60]
    assert oracle.features_per_level(2000).tolist() == [434, 362, 302, 251, 209, 175, 145, 122]
    assert oracle.features_per_level(4000).tolist() == [869, 724, 603, 503, 419, 349, 291, 242]
    assert int((w.astype(np.int64) * h).sum()) == 950532


def test_db_scores_and_topk(oracle):
    rng = np.random.default_rng(5)
    q = rng.integers(0, 256, (3, 40, 32), dtype=np.uint8)
    db = rng.integers(0, 256, (7, 50, 32), dtype=np.uint8)
    db[4, :40] = q[1]  # planted copy
    qc = np.array([40, 40, 33], np.int32)
    kc = np.array([50, 50, 10, 50, 50, 0, 50], np.int32)
    sc = oracle.db_scores(q, qc, db, kc)
    for i in range(3):
        for k in range(7):
            assert sc[i, k] == len(oracle.match_ratio(q[i, :qc[i]], db[k, :kc[k]])[0])
    assert sc[1, 4] >= 40
    s, ids = oracle.db_topk(sc, 3, id_base=100)
    assert ids[1, 0] == 104
    for i in range(3):
        order = sorted(range(7), key=lambda k: (-sc[i, k], k))[:3]
        assert ids[i].tolist() == [100 + k for k in order] and s[i].tolist() == [sc[i, k] for k in order]
