#!/usr/bin/env python
"""bench_sub.py -- the sub-records bench.py adds to its ONE JSON line at every N, so that the driver's BENCH / SCALE records
carry the second half of BASELINE.json's metric ("DB Hamming matches/s @1/2/4/8 GPU") and configs[2], not only configs[1]:

  "db": keyframe-database scoring (configs[4]) through the C ABI's sharded entry point -- every rank scores its resident shard
        on the tensor cores, keeps a local top-k, ONE ncclAllGather + merge inside the library (mslam_b200_db_query_sharded_dev),
        no host synchronisation in the step.  Reported: whole-job and per-GPU descriptor pairs/s, the roofline against the int8
        tensor peak MEASURED in the same run (mslam_b200_measure_tensor_peak), the cost of the collective + merge alone
        (an empty-shard query), a CPU baseline on a bounded sample, and a checksum of the merged top-k of a small COMMON
        database that every N shards differently -- equal on all ranks, equal to the CPU oracle's answer, hence equal across N.
  "c3": configs[2] (1280x720, 2000 keypoints, cross-checked matching), device-resident frames/s over a short run.

Both run AFTER the headline c2 measurement, outside its timed region.
"""
import hashlib
import os
import time

import numpy as np

DB_KF_PER_GPU = 12500
DB_DESC = 1000
DB_QUERY_BATCH = 64
DB_TOPK = 10
COMMON_KF, COMMON_Q = 384, 8


def _device_random_bytes(torch, shape, seed, device):
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    return torch.randint(0, 256, shape, dtype=torch.uint8, device=device, generator=g)


def _max_over_ranks(torch, dist, world, ms, dev):
    if world > 1:
        tt = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return float(tt.item())
    return ms


def db_subrecord(pkg, torch, dist, rank, world, local, stream, steps=3, warmup=2, cpu_seconds=6.0, want_cpu=True):
    from modular_slam_b200.keyframe_db import ShardedKeyframeDb, shard_range
    dev = torch.device("cuda", local)
    kf_local = int(os.environ.get("MSLAM_BENCH_KF_PER_GPU", DB_KF_PER_GPU))
    n, qb, k = DB_DESC, DB_QUERY_BATCH, DB_TOPK
    ctx = pkg.Context(device=local, max_width=64, max_height=64, max_batch=1, nfeatures=n, stream=stream.cuda_stream)
    db = ShardedKeyframeDb(ctx, rank, world)  # world > 1: NCCL id from rank 0 over torch.distributed, ncclCommInitRank in the library
    k0, k1 = shard_range(kf_local * world, rank, world)
    n_batches = 4
    d_queries = _device_random_bytes(torch, (n_batches * qb, n, 32), 2, dev)  # the same on every rank
    d_qc = torch.full((n_batches * qb,), n, dtype=torch.int32, device=dev)
    d_db = _device_random_bytes(torch, (k1 - k0, n, 32), 1 + 7919 * rank, dev)
    # plant: the first keyframe of every shard is a copy of query frame (rank % qb) of every batch position -> known top-1
    for bi in range(n_batches):
        d_db[bi] = d_queries[bi * qb + rank % qb]
    d_kc = torch.full((k1 - k0,), n, dtype=torch.int32, device=dev)
    db.load_shard(d_db, d_kc, k0)
    torch.cuda.synchronize()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(i):
        q0 = (i % n_batches) * qb
        return db.query(d_queries[q0:q0 + qb], d_qc[q0:q0 + qb], k)

    for i in range(warmup):
        s, ids = step(i)
    barrier()
    # planted copies: query frame r of batch b is keyframe (shard r's first id + b), score = all descriptors matched
    s_h, i_h = s.cpu().numpy(), ids.cpu().numpy()
    bi = (warmup - 1) % n_batches
    planted_ok = all(int(i_h[r % qb, 0]) == shard_range(kf_local * world, r, world)[0] + bi and int(s_h[r % qb, 0]) == n
                     for r in range(min(world, qb)))
    launches0 = ctx.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for i in range(steps):
        step(warmup + i)
    e1.record(stream)
    barrier()
    ms_total = _max_over_ranks(torch, dist, world, e0.elapsed_time(e1), dev)
    launches = ctx.launch_count - launches0
    # the scoring kernel alone (stage timing brackets it with events on the context's stream), rank 0's shard
    ctx.enable_stage_timing(True)
    for i in range(steps):
        step(warmup + i)
    st = ctx.stage_times()
    ctx.enable_stage_timing(False)
    db_ms = st["db"][0] / max(st["db"][1], 1)
    # the collective + merge alone: a query against an EMPTY shard is memset -> ncclAllGather -> merge
    os_ = torch.empty((qb, k), dtype=torch.int32, device=dev)
    oi = torch.empty((qb, k), dtype=torch.int32, device=dev)
    for _ in range(3):
        ctx.db_query_sharded_dev(d_queries[:qb], d_qc[:qb], qb, n, None, None, 0, n, 0, k, os_, oi)
    barrier()
    reps = 20
    e0.record(stream)
    for _ in range(reps):
        ctx.db_query_sharded_dev(d_queries[:qb], d_qc[:qb], qb, n, None, None, 0, n, 0, k, os_, oi)
    e1.record(stream)
    barrier()
    coll_ms = _max_over_ranks(torch, dist, world, e0.elapsed_time(e1) / reps, dev)
    # a small COMMON database (seeded numpy, the same for every N), sharded over the ranks of THIS run
    rng = np.random.default_rng(77)
    cq = rng.integers(0, 256, (COMMON_Q, n, 32), dtype=np.uint8)
    cdb = pkg.synth.descriptor_db(COMMON_KF, n, seed=78)
    pkg.synth.plant_copies(cdb, cq, frac=0.05, seed=79)
    c0, c1 = shard_range(COMMON_KF, rank, world)
    d_cdb = torch.from_numpy(cdb[c0:c1].copy()).to(dev)
    d_ckc = torch.full((c1 - c0,), n, dtype=torch.int32, device=dev)
    d_cq = torch.from_numpy(cq).to(dev)
    d_cqc = torch.full((COMMON_Q,), n, dtype=torch.int32, device=dev)
    cs = torch.empty((COMMON_Q, k), dtype=torch.int32, device=dev)
    ci = torch.empty((COMMON_Q, k), dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    ctx.db_query_sharded_dev(d_cq, d_cqc, COMMON_Q, n, d_cdb if c1 > c0 else None, d_ckc if c1 > c0 else None, c1 - c0, n, c0, k,
                             cs, ci)
    ctx.synchronize()
    common = np.stack([cs.cpu().numpy(), ci.cpu().numpy()])
    digest = hashlib.sha256(common.tobytes()).hexdigest()[:16]
    digests = [digest]
    if world > 1:
        digests = [None] * world
        dist.all_gather_object(digests, digest)
    barrier()
    db.close()  # every rank leaves the database communicator together, before any of them can exit
    barrier()
    if rank != 0:
        ctx.close()
        return None
    pairs_per_step = float(qb) * kf_local * world * n * n
    value = pairs_per_step * steps / (ms_total * 1e-3)
    per_gpu = float(qb) * kf_local * n * n / (db_ms * 1e-3)
    tp = ctx.measure_tensor_peak()
    peak = tp["int8_ops_per_s_n128_sustained"]
    rec = {"metric": "DB Hamming descriptor-pair evaluations/sec", "value": value, "unit": "descriptor pairs/s",
           "per_gpu_value": value / world, "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": ms_total / steps,
           "config": {"workload": "configs[4]: query frames x sharded keyframe-descriptor DB, kNN(2) + ratio 0.7 match count per "
                                  "(query frame, keyframe), per-query top-k, ncclAllGather + merge inside the C ABI",
                      "keyframes_per_gpu": kf_local, "keyframes_total": kf_local * world, "descriptors_per_frame": n,
                      "query_frames_per_step": qb, "top_k": k,
                      "entry_point": "mslam_b200_db_attach + mslam_b200_db_query_sharded_dev (one stream, no host synchronisation)"},
           "gpu_launches": int(launches),
           "roofline": {"bound": "tensor", "kernel": "k_knn2_mma_wide<true, 8> (+ k_expand_pm1)", "achieved": per_gpu * 512 / 1e12,
                        "peak": peak / 1e12, "unit": "TOP/s (int8; algorithmic: 2 x 256 per descriptor pair)",
                        "frac": per_gpu * 512 / peak, "traffic": None,
                        "peak_source": "tcgen05.mma.kind::i8 M128 N128 K32 issued back to back on all SMs from resident operands, "
                                       "sustained, measured in this run (mslam_b200_measure_tensor_peak)",
                        "peaks_measured": {k_: v / 1e12 for k_, v in tp.items()}, "kernel_ms_per_launch": db_ms,
                        "pairs_per_s_per_gpu_kernel_only": per_gpu},
           "collective": {"allgather_plus_merge_ms": coll_ms, "bytes_per_rank": 2 * qb * k * 4,
                          "share_of_step": coll_ms / (ms_total / steps),
                          "how": "an empty-shard query: memset -> ncclAllGather -> k_topk merge, max over ranks, mean of 20"},
           "checks": {"planted_copies_top1": bool(planted_ok), "common_db_digest": digest,
                      "common_db_digest_equal_on_all_ranks": all(d == digest for d in digests),
                      "common_db": f"{COMMON_Q} query frames x {COMMON_KF} keyframes x {n} descriptors (numpy seeds 77-79), sharded "
                                   f"over the {world} rank(s) of this run; the digest is the same for every N"}}
    if want_cpu:
        from oracle import oracle as O
        threads = os.cpu_count() or 1
        full = np.full
        t0 = time.perf_counter()
        want = O.db_topk(O.db_scores(cq, full(COMMON_Q, n, np.int32), cdb, full(COMMON_KF, n, np.int32), nthreads=threads), k)
        t_common = time.perf_counter() - t0
        rec["checks"]["common_db_equals_cpu_oracle"] = bool((common[0] == want[0]).all() and (common[1] == want[1]).all())
        nk = 8 * threads
        q1, db1 = cq[:1], cdb[:nk] if nk <= COMMON_KF else pkg.synth.descriptor_db(nk, n, seed=5)
        best, reps_done, t_spent = 1e9, 0, 0.0
        while t_spent < cpu_seconds and reps_done < 50:
            t0 = time.perf_counter()
            O.db_scores(q1, full(1, n, np.int32), db1, full(len(db1), n, np.int32), nthreads=threads)
            dt = time.perf_counter() - t0
            best, reps_done, t_spent = min(best, dt), reps_done + 1, t_spent + dt
        rec["cpu_baseline"] = {"value": float(len(db1)) * n * n / best, "unit": "descriptor pairs/s", "cores": threads, "kind": "port",
                               "sample": f"1 query frame x {len(db1)} keyframes x {n}^2 pairs, best of {reps_done}; OpenMP over keyframes "
                                         f"(the {COMMON_Q} x {COMMON_KF} common database took {t_common:.2f} s on the same cores)",
                               "extrapolation": "rate only: the 64 x 12,500 x N step is not run on the CPU"}
    ctx.close()
    return rec


def frames_subrecord(pkg, torch, dist, rank, world, local, stream, spec, steps=6, warmup=3, batch=None, want_cpu=True,
                     cpu_seconds=6.0):
    """device-resident frames/s of one of bench.WORKLOADS (spec = (W, H, nfeatures, batch, match mode, label)) over a short run,
    same step as the headline: detect the current frame of every sequence + match against the previous one"""
    W, H, nfeat, B, mode, label = spec
    B = batch or B
    dev = torch.device("cuda", local)
    T = 2
    bases, frames = None, []
    for t in range(T + 1):
        fr, bases = pkg.synth.frame_batch(W, H, B, t, seed=1234 + 100003 * rank, bases=bases)
        frames.append(fr)
    ctx = pkg.Context(device=local, max_width=W, max_height=H, max_batch=B, nfeatures=nfeat, stream=stream.cuda_stream)
    cap = ctx.kp_cap
    d_frames = [torch.from_numpy(f).to(dev) for f in frames]
    d_kp = [torch.zeros((B, cap, 6), dtype=torch.float32, device=dev) for _ in range(2)]
    d_desc = [torch.zeros((B, cap, 32), dtype=torch.uint8, device=dev) for _ in range(2)]
    d_cnt = [torch.zeros(B, dtype=torch.int32, device=dev) for _ in range(2)]
    d_m = [torch.zeros((B, cap), dtype=torch.int32, device=dev) for _ in range(3)]
    d_mn = torch.zeros(B, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()

    def step(t, cur):
        ctx.detect_batch_dev(d_frames[t], B, W, H, d_kp[cur], d_desc[cur], cap, d_cnt[cur])
        ctx.match_batch_dev(mode, d_desc[cur], d_cnt[cur], d_desc[cur ^ 1], d_cnt[cur ^ 1], B, cap, d_m[0], d_m[1], d_m[2], d_mn)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    ctx.detect_batch_dev(d_frames[0], B, W, H, d_kp[0], d_desc[0], cap, d_cnt[0])
    cur = 1
    for i in range(warmup):
        step(1 + i % T, cur)
        cur ^= 1
    barrier()
    n_kp, n_m = int(d_cnt[cur ^ 1].sum().item()), int(d_mn.sum().item())
    launches0 = ctx.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for i in range(steps):
        step(1 + i % T, cur)
        cur ^= 1
    e1.record(stream)
    barrier()
    ms = _max_over_ranks(torch, dist, world, e0.elapsed_time(e1), dev)
    launches = ctx.launch_count - launches0
    # per-stage times (second pass, one stream)
    ctx.set_flags(pkg.FLAG_SERIAL)
    ctx.enable_stage_timing(True)
    for i in range(steps):
        step(1 + i % T, cur)
        cur ^= 1
    stage = ctx.stage_times()
    ctx.enable_stage_timing(False)
    ctx.close()
    if rank != 0:
        return None
    w, h = None, None
    rec = {"metric": "ORB extract+match frames/sec", "value": B * steps * world / (ms * 1e-3), "unit": "frames/s", "n_gpus": world,
           "steps": steps, "warmup": warmup, "ms_per_step": ms / steps,
           "config": {"workload": label, "frame": [W, H], "nfeatures": nfeat, "nlevels": 8, "batch_per_gpu": B,
                      "match": "crosscheck" if mode else "knn2+ratio0.7"},
           "gpu_launches": int(launches),
           "stages_ms_per_step": {s: v[0] / steps for s, v in stage.items() if v[1] > 0},
           "checks": {"keypoints_per_batch": n_kp, "matches_per_batch": n_m}}
    if want_cpu:
        import bench
        from oracle import oracle as O
        threads = os.cpu_count() or 1
        ns = min(B, max(8, threads))
        best, reps_done, t_spent = 1e9, 0, 0.0
        while t_spent < cpu_seconds and reps_done < 10:
            dt = bench.cpu_port_time(pkg, O, frames[0][:ns], frames[1][:ns], nfeat, mode, threads)[0]
            best, reps_done, t_spent = min(best, dt), reps_done + 1, t_spent + dt
        rec["cpu_baseline"] = {"value": ns / best, "unit": "frames/s", "cores": threads, "kind": "port",
                               "sample": f"{ns} frames (detect + match vs previous frame), best of {reps_done}; OpenMP over frames"}
    return rec
