"""CPU, world_size 2 over gloo: the host logic of the multi-GPU path (sharding.py) -- row shards with global
ids, the single all-gather of [Q,k] candidate lists and the (distance, index) merge reproduce the search over
the whole database.  The per-shard search and the merge are the ORACLE here (no CUDA device in this test); on a
GPU box the same functions run with the library (tools/sharded_nn_check.py, tests marked gpu)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from multi_robot_slam_separators_b200 import sharding  # noqa: E402


def numpy_merge(all_idx, all_dist):
    """reference of sepb_topk_merge_dev: per row the k smallest (dist, idx) over all lists, -1/inf padded"""
    a_i, a_d = all_idx.numpy(), all_dist.numpy()
    lists, Q, k = a_i.shape
    out_i = np.full((Q, k), -1, np.int32)
    out_d = np.full((Q, k), np.inf)
    for q in range(Q):
        cand = [(a_d[l, q, j], a_i[l, q, j]) for l in range(lists) for j in range(k) if a_i[l, q, j] >= 0]
        cand.sort()
        for j, (d, i) in enumerate(cand[:k]):
            out_i[q, j], out_d[q, j] = i, d
    return torch.from_numpy(out_i), torch.from_numpy(out_d)


def _worker(rank, world, port, n_rows, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import find_matches as fm
    rng = np.random.default_rng(11)
    db = rng.standard_normal((n_rows, 16)) * 0.1
    if n_rows > 7:
        db[5] = db[2]                                  # exact tie across / inside shards -> lowest global id first
    q = rng.standard_normal((9, 16)) * 0.1
    q[0] = db[2] if n_rows > 2 else q[0]
    lo, hi = sharding.shard_rows(n_rows, rank, world)
    k = 3

    def local():
        idx, d = fm.topk(q, db[lo:hi], k) if hi > lo else (np.full((9, k), -1, np.int32), np.full((9, k), np.inf))
        idx = np.where(idx >= 0, idx + lo, -1).astype(np.int32)     # global ids, like sepb_db_set_base
        return torch.from_numpy(idx), torch.from_numpy(d)

    got_i, got_d = sharding.sharded_topk(local, numpy_merge)
    want_i, want_d = fm.topk(q, db, k)
    ok = np.array_equal(got_i.numpy(), want_i) and np.array_equal(got_d.numpy(), want_d)
    # pair blocks tile the candidate list exactly
    blocks = [sharding.pair_block(45, r, world) for r in range(world)]
    ok = ok and blocks[0][0] == 0 and blocks[-1][1] == 45 and all(blocks[i][1] == blocks[i + 1][0] for i in range(world - 1))
    t = torch.tensor([1 if ok else 0])
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if rank == 0:
        ret.put(int(t.item()))
    dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("n_rows", [40, 7, 1])
def test_sharded_topk_world2_gloo(n_rows):
    ctx = mp.get_context("spawn")
    ret = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_rows, ret)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert ret.get(timeout=5) == 1


def test_shard_rows_cover():
    for n in (0, 1, 7, 8, 350000):
        for w in (1, 2, 3, 8):
            r = [sharding.shard_rows(n, i, w) for i in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(r, r[1:]))
            assert max(h - l for l, h in r) - min(h - l for l, h in r) <= 1
    with pytest.raises(ValueError):
        sharding.shard_rows(5, 2, 2)
