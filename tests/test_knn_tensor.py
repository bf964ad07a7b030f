"""GPU parity tests of the tensor-core kNN / keyframe-DB scoring kernel (csrc/knn_mma.cu), forced on with
MSLAM_B200_FLAG_KNN_TENSOR and compared bit for bit with the CPU oracle (oracle/orb_oracle.c: orc_knn2 / orc_match_*,
orc_db_scores) and with the scalar XOR + POPC kernel (MSLAM_B200_FLAG_KNN_SCALAR).  The cases sit on the kernel's seams:
128-descriptor key windows and 256-descriptor train tiles (the N = 256 form: the two column halves of a tile are drained by
different warps, so last tiles that end inside the first half, on the half boundary and inside the second half all occur),
256-query CTAs, the ring and accumulator-stage wrap-around, ragged last tiles,
distance extremes (0 and 256) and ties (equal distance -> lower train index)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tensor_ctx(pkg):
    with pkg.Context(max_width=64, max_height=64, max_batch=8, nfeatures=100, max_descriptors=1400,
                     flags=pkg.FLAG_KNN_TENSOR) as c:
        yield c


@pytest.fixture(scope="module")
def scalar_ctx(pkg):
    with pkg.Context(max_width=64, max_height=64, max_batch=8, nfeatures=100, max_descriptors=1400,
                     flags=pkg.FLAG_KNN_SCALAR) as c:
        yield c


def _sets(rng, n_q, n_t, copies=0.3):
    q = rng.integers(0, 256, (n_q, 32), dtype=np.uint8)
    t = rng.integers(0, 256, (n_t, 32), dtype=np.uint8)
    m = int(min(n_q, n_t) * copies)
    if m:
        t[rng.permutation(n_t)[:m]] = q[rng.permutation(n_q)[:m]]  # true matches at random places
    return q, t


@pytest.mark.parametrize("n_q,n_t", [(1, 1), (1, 2), (5, 127), (127, 128), (128, 129), (129, 255), (255, 256), (256, 257),
                                     (257, 511), (300, 512), (513, 513), (640, 1025), (1400, 1400), (1400, 3), (3, 1400),
                                     (1279, 1281), (300, 383), (300, 384), (300, 385), (700, 641), (260, 897), (131, 769)])
def test_knn2_sizes_on_tile_seams(tensor_ctx, scalar_ctx, oracle, n_q, n_t):
    rng = np.random.default_rng(n_q * 7919 + n_t)
    q, t = _sets(rng, n_q, n_t)
    got = tensor_ctx.knn2(q, t)
    ref = oracle.knn2(q, t)
    sca = scalar_ctx.knn2(q, t)
    for g, r, s, what in zip(got, ref, sca, ("idx0", "dist0", "idx1", "dist1")):
        assert (np.asarray(g) == np.asarray(r)).all(), f"{what} vs oracle ({n_q} x {n_t})"
        assert (np.asarray(g) == np.asarray(s)).all(), f"{what} vs scalar kernel ({n_q} x {n_t})"


def test_ties_and_distance_extremes(tensor_ctx, oracle):
    rng = np.random.default_rng(3)
    base = rng.integers(0, 256, (1, 32), dtype=np.uint8)
    # every train descriptor identical: distance ties everywhere -> indices 0 and 1
    t = np.repeat(base, 700, axis=0)
    q = np.concatenate([base, ~base, rng.integers(0, 256, (298, 32), dtype=np.uint8)])  # distance 0, distance 256, random
    i0, d0, i1, d1 = tensor_ctx.knn2(q, t)
    assert (i0 == 0).all() and (i1 == 1).all() and d0[0] == 0 and d0[1] == 256 and (d0 == d1).all()
    r = oracle.knn2(q, t)
    assert all((np.asarray(a) == np.asarray(b)).all() for a, b in zip((i0, d0, i1, d1), r))
    # few distinct descriptors repeated: many ties at several distances, winners must be the lowest indices
    pool = rng.integers(0, 256, (4, 32), dtype=np.uint8)
    t = pool[rng.integers(0, 4, 900)]
    q = pool[rng.integers(0, 4, 500)]
    q[::3, 5] ^= 0x11
    got = tensor_ctx.knn2(q, t)
    ref = oracle.knn2(q, t)
    assert all((np.asarray(a) == np.asarray(b)).all() for a, b in zip(got, ref))
    # all bits differ from every train descriptor
    t = np.repeat(base, 130, axis=0)
    got = tensor_ctx.knn2(np.repeat(~base, 3, axis=0), t)
    assert (got[1] == 256).all() and (got[3] == 256).all() and (got[0] == 0).all() and (got[2] == 1).all()


def test_matchers_and_ragged_batches(pkg, tensor_ctx, oracle):
    rng = np.random.default_rng(11)
    for n_from, n_to in ((2, 1), (700, 900), (1314, 3), (513, 129), (1000, 1001)):
        f, t = _sets(rng, n_from, n_to, copies=0.5)
        got = tensor_ctx.match_knn_ratio(f, t)
        ref = oracle.match_ratio(f, t)
        assert all((np.asarray(g).astype(np.int64) == r).all() for g, r in zip(got, ref)), (n_from, n_to)
        got = tensor_ctx.match_crosscheck(t, f)
        ref = oracle.match_crosscheck(t, f)
        assert all((g == r).all() for g, r in zip(got, ref)), (n_from, n_to)
    # a ragged batch: per-pair counts from 0 to the capacity, one kernel launch for all pairs
    B, cap = 8, 1400
    nf = np.array([1400, 0, 1, 129, 1024, 257, 640, 2], np.int32)
    nt = np.array([1400, 300, 1, 0, 1023, 1400, 128, 1400], np.int32)
    frm = rng.integers(0, 256, (B, cap, 32), dtype=np.uint8)
    to = rng.integers(0, 256, (B, cap, 32), dtype=np.uint8)
    for b in range(B):
        m = min(nf[b], nt[b]) // 2
        to[b, :m] = frm[b, :m]
    a, bb, d, n = tensor_ctx.match_batch(0, frm, nf, to, nt)
    for b in range(B):
        fi, ti, di = oracle.match_ratio(frm[b, :nf[b]], to[b, :nt[b]])
        assert n[b] == len(fi), b
        assert (a[b, :n[b]] == fi).all() and (bb[b, :n[b]] == ti).all() and (d[b, :n[b]] == di).all(), b
    a, bb, d, n = tensor_ctx.match_batch(1, frm, nf, to, nt)  # cross-check: query = `to`, train = `from`
    for b in range(B):
        qi, tj, dj = oracle.match_crosscheck(to[b, :nt[b]], frm[b, :nf[b]])
        assert n[b] == len(qi), b
        assert (bb[b, :n[b]] == qi).all() and (a[b, :n[b]] == tj).all() and (d[b, :n[b]] == dj).all(), b


@pytest.mark.parametrize("ratio", [0.7, 0.8])
def test_db_scores_vs_oracle(pkg, oracle, ratio):
    rng = np.random.default_rng(5)
    q = rng.integers(0, 256, (5, 600, 32), dtype=np.uint8)
    db = pkg.synth.descriptor_db(70, 520, seed=9)
    pkg.synth.plant_copies(db, q[:, :520], frac=0.2, seed=4)
    qc = np.array([600, 599, 257, 2, 1], np.int32)
    kc = rng.integers(1, 521, 70).astype(np.int32)
    kc[3], kc[7], kc[9], kc[11] = 0, 33, 520, 256
    with pkg.Context(max_width=64, max_height=64, max_batch=8, nfeatures=100, max_descriptors=600, ratio=ratio,
                     flags=pkg.FLAG_KNN_TENSOR) as c:
        s, ids, allsc = c.db_query(q, qc, db, kc, k=5, id_base=1000, want_scores=True)
        again = c.db_query(q, qc, db, kc, k=5, id_base=1000, want_scores=True)[2]  # scores are re-zeroed per call
    if ratio == 0.7:
        ref = oracle.db_scores(q, qc, db, kc, nthreads=4)
    else:  # score(Q, K) = |OrbOpenCvMatcher::match(from = Q, to = K)| with the other ratio, pair by pair
        ref = np.array([[len(oracle.match_ratio(q[i, :qc[i]], db[k, :kc[k]], ratio)[0]) for k in range(len(kc))]
                        for i in range(len(qc))], np.int32)
    assert (allsc == ref).all() and (again == ref).all()
    rs, rid = oracle.db_topk(ref, 5, id_base=1000)
    assert (s == rs).all() and (ids == rid).all()
