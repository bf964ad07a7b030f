"""Sharded keyframe-descriptor database (BASELINE.json configs[4]).

The reference only sketches loop-candidate verification (rgbd_feature_frontend.cpp:499-533: for each candidate
keyframe run matchLandmarks, score = what survives); the build definition is

    score(Q, K) = | OrbOpenCvMatcher::match(from = Q, to = K) |        (orb_feature.cpp:84-117 semantics)
    per query frame: the k keyframes with the best (score desc, keyframe id asc)

Multi-GPU layout: keyframes are sharded by contiguous id range, one shard per rank, resident in that rank's HBM;
every rank sees all query frames; each rank scores its shard and keeps a local top-k; ONE all-gather of
[n_query_frames][k] (score, id) pairs per rank (NCCL over NVLink on GPUs) and a deterministic k-way merge with the
same ordering give every rank the global answer -- bit-identical to the single-GPU result.

The collective plumbing (`shard_range`, `exchange_topk`) is backend-agnostic so that it is covered by world_size-2
gloo tests on CPU; the scoring and the merge are CUDA kernels behind the C ABI (mslam_b200_db_*), no CPU fallback.
"""
import torch
import torch.distributed as dist


def shard_range(n_keyframes, rank, world):
    """contiguous id range [k0, k1) of `rank`: the first n % world ranks hold one keyframe more"""
    if world < 1 or not (0 <= rank < world) or n_keyframes < 0:
        raise ValueError("bad shard request")
    base, extra = divmod(n_keyframes, world)
    k0 = rank * base + min(rank, extra)
    return k0, k0 + base + (1 if rank < extra else 0)


def exchange_topk(local_score, local_id, group=None):
    """all-gather the per-rank [Q][k] int32 top-k lists -> ([world][Q][k] scores, [world][Q][k] ids), the layout
    mslam_b200_db_merge_topk_dev reads ([n_lists][n_query_frames][k]).  One collective: score and id travel stacked."""
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


class ShardedKeyframeDb:
    """one rank's shard of the keyframe database + the query path (device tensors in, device tensors out)"""

    def __init__(self, ctx, rank=0, world=1, group=None):
        self.ctx, self.rank, self.world, self.group = ctx, rank, world, group
        self.d_db = self.d_counts = None
        self.id_base = 0
        self.n_local = 0
        self._scratch = {}

    def load_shard(self, d_db, d_counts, id_base):
        """d_db: uint8 CUDA tensor [n_local][cap][32]; d_counts: int32 CUDA tensor [n_local]; ids = id_base + row"""
        if d_db.dtype != torch.uint8 or d_db.dim() != 3 or d_db.shape[2] != 32 or not d_db.is_cuda:
            raise ValueError("shard must be a CUDA uint8 tensor [n][cap][32]")
        self.d_db, self.d_counts, self.id_base, self.n_local = d_db.contiguous(), d_counts.contiguous(), int(id_base), d_db.shape[0]

    def _buf(self, name, shape):
        t = self._scratch.get(name)
        if t is None or tuple(t.shape) != tuple(shape):
            t = torch.empty(shape, dtype=torch.int32, device=self.d_db.device)
            self._scratch[name] = t
        return t

    def score_local(self, d_queries, d_q_counts):
        """-> int32 [Q][n_local] match counts of every query frame against every keyframe of this shard"""
        Q, qcap, _ = d_queries.shape
        sc = self._buf("scores", (Q, self.n_local))
        self.ctx.db_scores_dev(d_queries, d_q_counts, Q, qcap, self.d_db, self.d_counts, self.n_local, self.d_db.shape[1], sc)
        return sc

    def query(self, d_queries, d_q_counts, k):
        """-> (scores [Q][k], ids [Q][k]) int32 CUDA tensors, identical on every rank.  Asynchronous on the context's
        stream up to the collective; the collective and the merge are ordered after it."""
        Q = d_queries.shape[0]
        sc = self.score_local(d_queries, d_q_counts)
        ls, li = self._buf("ls", (Q, k)), self._buf("li", (Q, k))
        self.ctx.db_topk_dev(sc, Q, self.n_local, self.id_base, k, ls, li)
        if self.world == 1:
            return ls, li
        self.ctx.synchronize()  # the context runs on its own stream; NCCL on torch's
        gs, gi = exchange_topk(ls, li, self.group)
        torch.cuda.current_stream().synchronize()
        os_, oi = self._buf("os", (Q, k)), self._buf("oi", (Q, k))
        self.ctx.db_merge_topk_dev(gs, gi, self.world, Q, k, os_, oi)
        return os_, oi
