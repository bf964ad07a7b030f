"""Sharded keyframe-descriptor database: the Python mirror of host/sharded_keyframe_db_b200.hpp (SURVEY.md section 8, rows
a8 and (e); BASELINE.json configs[4]), used by tests/ and bench.py.  The product path is the C ABI:

    mslam_b200_db_attach(ctx, rank, world, nccl_unique_id)       join the database's NCCL communicator
    mslam_b200_db_query_sharded_dev(ctx, queries, shard, k, out) score the shard -> local top-k -> ncclAllGather -> merge,
                                                                 one stream, no host synchronisation

The reference only sketches loop-candidate verification (rgbd_feature_frontend.cpp:499-533: for each candidate keyframe
run matchLandmarks, score = what survives); the build definition is

    score(Q, K) = | OrbOpenCvMatcher::match(from = Q, to = K) |        (orb_feature.cpp:84-117 semantics)
    per query frame: the k keyframes with the best (score desc, keyframe id asc)

Keyframes are sharded by contiguous id range, one shard per rank, resident in that rank's HBM; every rank sees all query
frames and ends with the same global top-k, bit-identical to the single-GPU result.

torch is used here for two things only: device tensors as buffers, and torch.distributed (any backend) to hand rank 0's NCCL
unique id to the other ranks.  `exchange_topk` restates the collective's layout over torch.distributed so that the host
logic is covered by world_size-2 gloo tests on CPU (tests/test_sharded_db.py); the GPU path does not call it.
"""
import numpy as np
import torch
import torch.distributed as dist

from .capi import NCCL_ID_BYTES, db_shard_range, nccl_unique_id


def shard_range(n_keyframes, rank, world):
    """contiguous id range [k0, k1) of `rank`: the first n % world ranks hold one keyframe more (the C ABI's rule)"""
    return db_shard_range(n_keyframes, rank, world)


def exchange_topk(local_score, local_id, group=None):
    """CPU-testable statement of the collective inside mslam_b200_db_query_sharded_dev: every rank contributes its
    [2][Q][k] (score, id) block, everyone receives [world][2][Q][k].  -> ([world][Q][k] scores, [world][Q][k] ids)"""
    if local_score.shape != local_id.shape or local_score.dtype != torch.int32 or local_id.dtype != torch.int32:
        raise ValueError("top-k lists must be int32 tensors of one shape")
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return local_score.unsqueeze(0).contiguous(), local_id.unsqueeze(0).contiguous()
    mine = torch.stack((local_score, local_id)).contiguous()          # [2][Q][k]
    gathered = torch.empty((world * 2,) + tuple(mine.shape[1:]), dtype=mine.dtype, device=mine.device)
    dist.all_gather_into_tensor(gathered, mine, group=group)          # concatenated along dim 0 (nccl and gloo agree)
    gathered = gathered.view((world, 2) + tuple(mine.shape[1:]))      # [world][2][Q][k]
    return gathered[:, 0].contiguous(), gathered[:, 1].contiguous()


def broadcast_unique_id(rank, group=None, device=None):
    """rank 0 asks the library for an NCCL unique id; torch.distributed (gloo or nccl) carries the 128 bytes to everyone"""
    buf = torch.zeros(NCCL_ID_BYTES, dtype=torch.uint8)
    if rank == 0:
        buf = torch.from_numpy(np.frombuffer(nccl_unique_id(), np.uint8).copy())
    if device is not None:
        buf = buf.to(device)
    dist.broadcast(buf, src=0, group=group)
    return buf.cpu().numpy().tobytes()


class ShardedKeyframeDb:
    """one rank's shard of the keyframe database + the query path (device tensors in, device tensors out).
    The context must run on the stream the caller orders its tensors on: pass
    Context(stream=torch.cuda.current_stream().cuda_stream), or synchronise around the calls."""

    def __init__(self, ctx, rank=0, world=1, group=None, unique_id=None):
        self.ctx, self.rank, self.world = ctx, rank, world
        self.d_db = self.d_counts = None
        self.id_base = 0
        self.n_local = 0
        self.k_cap = 1
        self._out = {}
        if world > 1:
            if unique_id is None:
                dev = "cuda" if dist.get_backend(group) == "nccl" else None
                unique_id = broadcast_unique_id(rank, group, dev)
            ctx.db_attach(rank, world, unique_id)

    def close(self):
        """leave the communicator (call on every rank once the last query has completed)"""
        self.ctx.synchronize()
        self.ctx.db_detach()

    def load_shard(self, d_db, d_counts, id_base):
        """d_db: uint8 CUDA tensor [n_local][cap][32] (n_local may be 0); d_counts: int32 CUDA tensor [n_local]"""
        if d_db.dtype != torch.uint8 or d_db.dim() != 3 or d_db.shape[2] != 32 or not d_db.is_cuda:
            raise ValueError("shard must be a CUDA uint8 tensor [n][cap][32]")
        self.d_db, self.d_counts, self.id_base, self.n_local = d_db.contiguous(), d_counts.contiguous(), int(id_base), d_db.shape[0]
        self.k_cap = max(int(d_db.shape[1]), 1)

    def score_local(self, d_queries, d_q_counts):
        """-> int32 [Q][n_local] match counts of every query frame against every keyframe of this shard (measurement aid)"""
        Q, qcap, _ = d_queries.shape
        sc = torch.empty((Q, self.n_local), dtype=torch.int32, device=d_queries.device)
        self.ctx.db_scores_dev(d_queries, d_q_counts, Q, qcap, self.d_db, self.d_counts, self.n_local, self.k_cap, sc)
        return sc

    def query(self, d_queries, d_q_counts, k):
        """-> (scores [Q][k], ids [Q][k]) int32 CUDA tensors, identical on every rank; asynchronous on the context's stream"""
        Q, qcap, _ = d_queries.shape
        key = (Q, k)
        if key not in self._out:
            self._out[key] = (torch.empty((Q, k), dtype=torch.int32, device=d_queries.device),
                              torch.empty((Q, k), dtype=torch.int32, device=d_queries.device))
        os_, oi = self._out[key]
        self.ctx.db_query_sharded_dev(d_queries, d_q_counts, Q, qcap, self.d_db if self.n_local else None,
                                      self.d_counts if self.n_local else None, self.n_local, self.k_cap, self.id_base, k, os_, oi)
        return os_, oi
