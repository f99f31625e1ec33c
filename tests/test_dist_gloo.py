"""World-size-2 `gloo` run on CPU of the sharded keyframe-database search protocol (BASELINE.json configs[4]):
rank r searches its shard with global indices (index_base), ranks all_gather their nq x 8-byte best-2 records, every rank
folds them. On CPU the per-shard search is the oracle and the fold is a numpy restatement of k_allpairs_merge; the result
must equal the oracle over the whole database, including ties across shard boundaries. The same protocol with the CUDA
kernels and NCCL is what bench.py runs at --gpus N."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def fold_best2(parts):
    """parts [n_parts, nq] of (idx u32, d1 u16, d2 u16) -> merged [nq]; mirrors csrc/match_allpairs.cu k_allpairs_merge."""
    n_parts, nq = parts.shape
    bd = np.full(nq, 256, np.int64); bi = np.full(nq, 0xFFFFFFFF, np.int64); sd = np.full(nq, 256, np.int64)
    for p in range(n_parts):
        d1, d2, ix = parts["d1"][p].astype(np.int64), parts["d2"][p].astype(np.int64), parts["idx"][p].astype(np.int64)
        better = (d1 < bd) | ((d1 == bd) & (ix < bi))
        sd = np.where(better, np.minimum(sd, bd), np.minimum(sd, d1))
        bd = np.where(better, d1, bd); bi = np.where(better, ix, bi)
        sd = np.minimum(sd, d2)
    return bi, bd, sd


def _worker(rank, world, port, q_path, t_path, out_path):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist
    from _oracle import OracleMatcher
    from motioncheck_ccm_slam_b200.matcher import BEST2_DTYPE
    from motioncheck_ccm_slam_b200.sharding import shard_range
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    q, t = np.load(q_path), np.load(t_path)
    lo, hi = shard_range(len(t), rank, world)
    res = OracleMatcher(0.75).allpairs_best2(q, t[lo:hi])
    part = np.zeros(len(q), BEST2_DTYPE)
    part["idx"] = np.where(res["best_idx"] >= 0, res["best_idx"] + lo, 0xFFFFFFFF)
    part["d1"], part["d2"] = res["best_dist"], res["best_dist2"]
    mine = torch.from_numpy(part.view(np.uint8).reshape(len(q), 8).copy())
    gathered = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(gathered, mine)
    parts = np.stack([g.numpy().view(BEST2_DTYPE).reshape(-1) for g in gathered])
    bi, bd, sd = fold_best2(parts)
    np.save(out_path % rank, np.stack([bi, bd, sd]))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_db_search_two_ranks_gloo(tmp_path):
    torch = pytest.importorskip("torch")
    import torch.multiprocessing as mp
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from _oracle import OracleMatcher
    rng = np.random.default_rng(11)
    nq, nt = 200, 6001
    t = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
    q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    q[:50] = t[rng.integers(0, nt, 50)]
    t[nt - 1] = t[5]; q[0] = t[5]          # duplicate target in the OTHER shard: lowest global index must win
    t[3001] = t[2999]; q[1] = t[2999]      # tie straddling the shard boundary (3000)
    qp, tp = str(tmp_path / "q.npy"), str(tmp_path / "t.npy")
    np.save(qp, q); np.save(tp, t)
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    out = str(tmp_path / "out%d.npy")
    mp.spawn(_worker, args=(2, port, qp, tp, out), nprocs=2, join=True)
    ref = OracleMatcher(0.75).allpairs_best2(q, t)
    for r in range(2):
        bi, bd, sd = np.load(out % r)
        assert np.array_equal(bi, ref["best_idx"]) and np.array_equal(bd, ref["best_dist"]) and np.array_equal(sd, ref["best_dist2"])
    assert ref["best_idx"][0] == 5 and ref["best_dist2"][0] == 0 and ref["best_idx"][1] == 2999


def test_partition_arithmetic():
    from motioncheck_ccm_slam_b200.sharding import agent_device, keyframe_shards, shard_range
    for n, w in [(4000000, 8), (7, 3), (0, 2), (5, 8)]:
        edges = [shard_range(n, r, w) for r in range(w)]
        assert edges[0][0] == 0 and edges[-1][1] == n
        assert all(edges[i][1] == edges[i + 1][0] for i in range(w - 1))
        assert max(h - l for l, h in edges) - min(h - l for l, h in edges) <= 1
    assert [agent_device(a, 2) for a in range(4)] == [0, 1, 0, 1]
    sh = keyframe_shards([1000] * 4000, 8)
    assert [s[1] - s[0] for s in sh] == [500] * 8 and sh[3][2] == 1500000 and sh[-1][3] == 4000000
    sh = keyframe_shards([900, 1100, 1000, 50, 950], 2)
    assert sh[0][0] == 0 and sh[-1][1] == 5 and sh[0][1] == sh[1][0] and sh[0][3] == sh[1][2] and sh[-1][3] == 4000
