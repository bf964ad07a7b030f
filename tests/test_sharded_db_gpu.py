"""GPU tests of the sharded keyframe database through the C ABI (mslam_b200_db_attach / _db_query_sharded_dev):
* a real shard shape (2,500 keyframes x 1,000 descriptors, 64 query frames x 1,000) -- the 8-query-frames-per-keyframe
  walk, the keyframe chunking and the ratio-count epilogue of k_knn2_mma<true> -- against oracle scores on a sampled subset,
  planted near-copies, and an exact top-k of the device's own score matrix;
* two GPUs in one process (one context per device, one thread per rank: NCCL all-gather inside the library), uneven and
  empty shards, == the single-GPU answer;  [needs 2 visible GPUs: `gpurun --gpus 2`; skipped otherwise]
* a tensor-core match on device 0 AND device 1 of one process (the dynamic shared-memory opt-in is per device).
"""
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _topk_numpy(scores, k, id_base=0):
    Q, K = scores.shape
    os_, oi = np.full((Q, k), -1, np.int32), np.full((Q, k), -1, np.int32)
    for q in range(Q):
        order = sorted(range(K), key=lambda j: (-int(scores[q, j]), j))[:k]
        os_[q, :len(order)] = scores[q, order]
        oi[q, :len(order)] = np.array(order, np.int32) + id_base
    return os_, oi


def test_real_shard_shape_through_the_sharded_entry_point(pkg, oracle):
    import torch
    K, n, Q, k = 2500, 1000, 64, 10
    queries = pkg.synth.descriptor_db(Q, n, seed=2)
    db = pkg.synth.descriptor_db(K, n, seed=1)
    planted = pkg.synth.plant_copies(db, queries, frac=0.01, seed=3)
    rng = np.random.default_rng(4)
    qc = np.full(Q, n, np.int32)
    kc = np.full(K, n, np.int32)
    qc[5], qc[17], qc[63] = 999, 513, 2          # ragged query frames
    ragged = rng.choice(K, 60, replace=False)
    kc[ragged] = rng.integers(0, n + 1, 60)      # ragged keyframes, some empty / single-descriptor
    kc[ragged[0]], kc[ragged[1]] = 0, 1
    with pkg.Context(max_width=64, max_height=64, max_batch=1, nfeatures=100, max_descriptors=n,
                     stream=torch.cuda.current_stream().cuda_stream) as ctx:
        d_q, d_qc = torch.from_numpy(queries).cuda(), torch.from_numpy(qc).cuda()
        d_db, d_kc = torch.from_numpy(db).cuda(), torch.from_numpy(kc).cuda()
        d_sc = torch.zeros((Q, K), dtype=torch.int32, device="cuda")
        os_ = torch.zeros((Q, k), dtype=torch.int32, device="cuda")
        oi = torch.zeros((Q, k), dtype=torch.int32, device="cuda")
        ctx.db_scores_dev(d_q, d_qc, Q, n, d_db, d_kc, K, n, d_sc)
        ctx.db_query_sharded_dev(d_q, d_qc, Q, n, d_db, d_kc, K, n, 40000, k, os_, oi)   # world 1: no attach needed
        ctx.synchronize()
        sc, s, ids = d_sc.cpu().numpy(), os_.cpu().numpy(), oi.cpu().numpy()
        # the integer-pipe kernel on the same shard: same score matrix
        ctx.set_flags(pkg.FLAG_KNN_SCALAR)
        d_sc2 = torch.zeros_like(d_sc)
        ctx.db_scores_dev(d_q, d_qc, Q, n, d_db, d_kc, K, n, d_sc2)
        ctx.synchronize()
        assert (d_sc2.cpu().numpy() == sc).all(), "tensor-core and integer-pipe score matrices differ"
    # oracle on a sample: planted pairs, the ragged keyframes, random pairs
    pairs = [(qi, kf) for kf, qi in planted.items()][:12]
    pairs += [(int(rng.integers(0, Q)), int(kf)) for kf in ragged[:12]]
    pairs += [(int(q), int(kf)) for q, kf in zip(rng.integers(0, Q, 16), rng.integers(0, K, 16))]
    pairs += [(5, 7), (17, 8), (63, 9)]
    for qi, kf in pairs:
        want = len(oracle.match_ratio(queries[qi, :qc[qi]], db[kf, :kc[kf]])[0])
        assert sc[qi, kf] == want, (qi, kf, sc[qi, kf], want)
    for kf, qi in planted.items():
        if qc[qi] == n and kc[kf] == n:
            assert sc[qi, kf] > n // 2 and ids[qi, 0] in [40000 + f for f, q2 in planted.items() if q2 == qi]
    ws, wi = _topk_numpy(sc, k, id_base=40000)
    assert (s == ws).all() and (ids == wi).all()


def _two_gpus():
    import torch
    return torch.cuda.device_count() >= 2


def test_tensor_core_match_on_two_devices_of_one_process(pkg, oracle, bundled):
    if not _two_gpus():
        pytest.skip("needs 2 GPUs")
    d0, d1 = bundled["desc0"], bundled["desc1"]
    want = oracle.match_ratio(d0, d1)
    for dev in (0, 1, 0):
        with pkg.Context(device=dev, max_width=64, max_height=64, max_batch=1, nfeatures=1000, max_descriptors=1000,
                         flags=pkg.FLAG_KNN_TENSOR) as ctx:
            fi, ti, di = ctx.match_knn_ratio(d0, d1)
            assert (fi.astype(np.int64) == want[0]).all() and (ti.astype(np.int64) == want[1]).all() and (di == want[2]).all()
            rng = np.random.default_rng(dev)
            q = rng.integers(0, 256, (3, 1600, 32), dtype=np.uint8)   # 51 KB of query frame: beyond the 48 KB default
            db = rng.integers(0, 256, (9, 300, 32), dtype=np.uint8)
            ctx.set_flags(pkg.FLAG_KNN_SCALAR)
            _, _, sc = ctx.db_query(q, np.full(3, 1600, np.int32), db, np.full(9, 300, np.int32), k=2, want_scores=True)
        assert (sc == oracle.db_scores(q, np.full(3, 1600, np.int32), db, np.full(9, 300, np.int32), nthreads=4)).all()


@pytest.mark.parametrize("K,k", [(37, 5), (1, 3), (2, 4)])
def test_two_ranks_in_one_process_equal_the_single_gpu_answer(pkg, oracle, K, k):
    """rank r = device r, one thread each; K = 1 leaves rank 1 with an empty shard, which must still join the collective"""
    if not _two_gpus():
        pytest.skip("needs 2 GPUs")
    import torch
    Q, n, world = 6, 256, 2
    rng = np.random.default_rng(K)
    queries = rng.integers(0, 256, (Q, n, 32), dtype=np.uint8)
    db = pkg.synth.descriptor_db(K, n, seed=K + 1)
    pkg.synth.plant_copies(db, queries, frac=0.3, seed=2)
    qc, kc = np.full(Q, n, np.int32), np.full(K, n, np.int32)
    want_sc = oracle.db_scores(queries, qc, db, kc, nthreads=4)
    ws, wi = oracle.db_topk(want_sc, k)
    uid = pkg.nccl_unique_id()
    results, errors = [None] * world, []

    def run(rank):
        try:
            torch.cuda.set_device(rank)
            with pkg.Context(device=rank, max_width=64, max_height=64, max_batch=1, nfeatures=100, max_descriptors=n,
                             flags=pkg.FLAG_KNN_TENSOR) as ctx:
                ctx.db_attach(rank, world, uid)
                assert ctx.db_shard_info() == (rank, world)
                k0, k1 = pkg.db_shard_range(K, rank, world)
                dev = f"cuda:{rank}"
                d_q, d_qc = torch.from_numpy(queries).to(dev), torch.from_numpy(qc).to(dev)
                d_db = torch.from_numpy(db[k0:k1].copy()).to(dev) if k1 > k0 else None
                d_kc = torch.from_numpy(kc[k0:k1].copy()).to(dev) if k1 > k0 else None
                os_ = torch.zeros((Q, k), dtype=torch.int32, device=dev)
                oi = torch.zeros((Q, k), dtype=torch.int32, device=dev)
                torch.cuda.synchronize(rank)
                for _ in range(3):  # repeated queries reuse the scratch and the communicator
                    ctx.db_query_sharded_dev(d_q, d_qc, Q, n, d_db, d_kc, k1 - k0, n, k0, k, os_, oi)
                ctx.synchronize()
                results[rank] = (os_.cpu().numpy(), oi.cpu().numpy())
                # host-in / host-out form of the same call
                hs, hi = ctx.db_query_sharded(queries, qc, d_db, d_kc, k1 - k0, n, k0, k)
                assert (hs == results[rank][0]).all() and (hi == results[rank][1]).all()
                ctx.db_detach()
        except Exception as e:  # noqa: BLE001
            errors.append((rank, repr(e)))

    threads = [threading.Thread(target=run, args=(r,)) for r in range(world)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=300)
    assert not errors, errors
    for r in range(world):
        assert results[r] is not None, f"rank {r} did not finish"
        assert (results[r][0] == ws).all() and (results[r][1] == wi).all(), (r, results[r], ws, wi)


def test_attach_argument_checks(pkg):
    with pkg.Context(max_width=64, max_height=64, max_batch=1, nfeatures=100) as ctx:
        assert ctx.db_shard_info() == (0, 1)
        ctx.db_attach(0, 1)  # world 1: no id, no NCCL
        with pytest.raises(pkg.MslamB200Error) as e:
            ctx.db_attach(2, 2, pkg.nccl_unique_id())
        assert e.value.status == pkg.E_INVALID
        with pytest.raises(pkg.MslamB200Error) as e:
            ctx.db_attach(0, 2, None)
        assert e.value.status == pkg.E_INVALID
        with pytest.raises(pkg.MslamB200Error) as e:
            ctx.set_flags(pkg.FLAG_KNN_SCALAR | pkg.FLAG_KNN_TENSOR)
        assert e.value.status == pkg.E_INVALID


# ---- the C++ class (host/sharded_keyframe_db_b200.hpp) through plugin_host_test ---------------------------------------------------
def _cpp_db_case(pkg, oracle, tmp_path, K=29, n=256, Q=5, k=4):
    import os
    rng = np.random.default_rng(K + Q)
    queries = rng.integers(0, 256, (Q, n, 32), dtype=np.uint8)
    db = pkg.synth.descriptor_db(K, n, seed=K)
    pkg.synth.plant_copies(db, queries, frac=0.3, seed=1)
    inp = tmp_path / "db_in.bin"
    with open(inp, "wb") as f:
        f.write(np.array([K, n, Q, k], np.int32).tobytes())
        f.write(db.tobytes())
        f.write(queries.tobytes())
    full = np.full
    ws, wi = oracle.db_topk(oracle.db_scores(queries, full(Q, n, np.int32), db, full(K, n, np.int32), nthreads=4), k)
    import __graft_entry__ as graft
    return str(inp), ws, wi, os.path.join(graft.PKG_DIR, "lib", "plugin_host_test")


def _read_cpp_db_out(path, Q, k):
    rec = np.fromfile(path, np.dtype([("id", "<i8"), ("score", "<i4"), ("pad", "<i4")])).reshape(Q, k)
    return rec["score"], rec["id"]


def test_cpp_sharded_db_class_one_rank(pkg, oracle, tmp_path):
    import subprocess
    inp, ws, wi, driver = _cpp_db_case(pkg, oracle, tmp_path)
    out = subprocess.run([driver, "db", inp, str(tmp_path / "o.bin")], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    s, ids = _read_cpp_db_out(tmp_path / "o.bin", *ws.shape)
    assert (s == ws).all() and (ids == wi).all()


def test_cpp_sharded_db_class_two_processes(pkg, oracle, tmp_path):
    """two C++ processes, one per GPU, NCCL id handed over through a file: both print the single-GPU answer"""
    if not _two_gpus():
        pytest.skip("needs 2 GPUs")
    import subprocess
    inp, ws, wi, driver = _cpp_db_case(pkg, oracle, tmp_path, K=31, n=300, Q=7, k=5)
    idfile = str(tmp_path / "nccl.id")
    procs = [subprocess.Popen([driver, "db", inp, str(tmp_path / f"o{r}.bin"), str(r), "2", idfile], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    for r, p in enumerate(procs):
        log, _ = p.communicate(timeout=300)
        assert p.returncode == 0, (r, log)
    for r in range(2):
        s, ids = _read_cpp_db_out(tmp_path / f"o{r}.bin", *ws.shape)
        assert (s == ws).all() and (ids == wi).all(), r
