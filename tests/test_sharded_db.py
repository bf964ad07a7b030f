"""Host-side logic of the sharded keyframe database: shard ranges, the all-gather layout and the
merge order, covered with world_size-2 gloo process groups on CPU.  The per-shard scores come from the CPU oracle here
(test infrastructure); on the GPU box tests/test_gpu_parity.py runs the same exchange over the CUDA kernels."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _merge_numpy(gs, gi, k):
    """[world][Q][k] lists -> top-k by (score desc, id asc), ignoring (-1, -1) pads"""
    W, Q, _ = gs.shape
    os_, oi = np.full((Q, k), -1, np.int32), np.full((Q, k), -1, np.int32)
    for q in range(Q):
        ent = [(int(s), int(i)) for s, i in zip(gs[:, q].ravel(), gi[:, q].ravel()) if s >= 0 and i >= 0]
        ent.sort(key=lambda e: (-e[0], e[1]))
        for r, (s, i) in enumerate(ent[:k]):
            os_[q, r], oi[q, r] = s, i
    return os_, oi


def _worker(rank, world, port, K, k, out_dir):
    sys.path.insert(0, ROOT)
    import __graft_entry__ as graft
    pkg = graft.load_package()
    from modular_slam_b200.keyframe_db import exchange_topk, shard_range
    from oracle import oracle as O
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(11)
        Q, n = 5, 96
        q = rng.integers(0, 256, (Q, n, 32), dtype=np.uint8)
        db = pkg.synth.descriptor_db(K, n, seed=21)
        pkg.synth.plant_copies(db, q, frac=0.3, seed=5)
        qc, kc = np.full(Q, n, np.int32), np.full(K, n, np.int32)
        if K > 1:
            kc[1] = 0
        k0, k1 = shard_range(K, rank, world)
        local = O.db_scores(q, qc, db[k0:k1], kc[k0:k1]) if k1 > k0 else np.zeros((Q, 0), np.int32)
        if k1 > k0:
            ls, li = O.db_topk(local, k, id_base=k0)
        else:
            ls, li = np.full((Q, k), -1, np.int32), np.full((Q, k), -1, np.int32)
        gs, gi = exchange_topk(torch.from_numpy(ls), torch.from_numpy(li))
        assert gs.shape == (world, Q, k) and gi.shape == (world, Q, k)
        assert (gs[rank].numpy() == ls).all() and (gi[rank].numpy() == li).all()  # layout: [rank][query][k]
        ms, mi = _merge_numpy(gs.numpy(), gi.numpy(), k)
        ws, wi = O.db_topk(O.db_scores(q, qc, db, kc), k)
        assert (ms == ws).all() and (mi == wi).all(), (rank, ms, ws, mi, wi)
        # every rank ends with the same answer
        mine = torch.from_numpy(np.stack([ms, mi]))
        other = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(other, mine)
        assert all((o == mine).all() for o in other)
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


def test_shard_ranges_partition_the_id_space():
    sys.path.insert(0, ROOT)
    import __graft_entry__ as graft
    graft.load_package()
    from modular_slam_b200.keyframe_db import shard_range
    for K in (0, 1, 7, 8, 100000, 12501):
        for world in (1, 2, 3, 8):
            rs = [shard_range(K, r, world) for r in range(world)]
            assert rs[0][0] == 0 and rs[-1][1] == K
            assert all(rs[i][1] == rs[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in rs]
            assert max(sizes) - min(sizes) <= 1
    assert shard_range(100000, 3, 8) == (37500, 50000)
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def test_exchange_is_identity_for_one_rank():
    sys.path.insert(0, ROOT)
    import __graft_entry__ as graft
    graft.load_package()
    from modular_slam_b200.keyframe_db import exchange_topk
    s = torch.arange(12, dtype=torch.int32).reshape(3, 4)
    gs, gi = exchange_topk(s, s + 100)
    assert gs.shape == (1, 3, 4) and (gs[0] == s).all() and (gi[0] == s + 100).all()
    with pytest.raises(ValueError):
        exchange_topk(s, (s + 1).to(torch.int64))


@pytest.mark.parametrize("K,k", [(23, 4), (2, 3), (1, 2)])
def test_sharded_topk_over_gloo_world2_equals_whole_db(tmp_path, K, k):
    """two ranks, uneven shards (and a rank with an empty shard when K = 1): merged result == whole-DB top-k"""
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), K, k, str(tmp_path)), nprocs=world, join=True)
    assert all(os.path.exists(tmp_path / f"ok{r}") for r in range(world))
