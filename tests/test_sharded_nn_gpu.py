"""Row-sharded NetVLAD database (GPU, through the C ABI): shards with global row ids + the device merge kernel and the
NCCL group give exactly the search over one database (ids and f64 distances bit-identical to scipy.cdist, reference
semantics R/scripts/data_handler.py:166-191).  One device is enough for the shard + merge logic; the NCCL test needs 2."""
import ctypes as C

import numpy as np
import pytest
import torch

from multi_robot_slam_separators_b200 import capi, sharding, synth
from oracle import find_matches as fm

pytestmark = pytest.mark.gpu


def make_db(n, dim, seed, nq):
    places = np.arange(n, dtype=np.int64)
    db = synth.make_netvlad(places, dim, seed)
    rng = np.random.default_rng(seed + 1)
    q = synth.make_netvlad(rng.integers(0, n, nq), dim, seed, noise=0.4)       # noisy revisits of database places
    db[n // 2 + 3] = db[7]                # an exact duplicate in another shard: ties resolve to the lowest GLOBAL id
    q[0] = db[7]
    return db, q


@pytest.mark.parametrize("mode", [capi.NN_EXACT, capi.NN_TENSOR])
def test_two_shards_on_one_device_merge_equals_one_database(mode):
    n, dim, nq, k = 3001, 128, 257, 5
    db, q = make_db(n, dim, 90, nq)
    ctx = capi.Context(0)
    try:
        whole = capi.NetvladDB(ctx, dim, n)
        whole.append(db)
        want_i, want_d = whole.topk(q, k, mode)
        ref_i, ref_d = fm.topk(q, db, k)
        assert np.array_equal(want_i, ref_i) and np.array_equal(want_d, ref_d)
        lists_i, lists_d = [], []
        for r in range(3):                                          # uneven shards (1001 / 1000 / 1000 rows)
            lo, hi = sharding.shard_rows(n, r, 3)
            s = capi.NetvladDB(ctx, dim, hi - lo, base=lo)
            s.append(db[lo:hi])
            i, d = s.topk(q, k, mode)
            assert i.min() >= lo and i.max() < hi                   # global ids
            lists_i.append(i); lists_d.append(d)
        all_i = torch.from_numpy(np.stack(lists_i)).cuda()
        all_d = torch.from_numpy(np.stack(lists_d)).cuda()
        got_i, got_d = sharding.merge_on_device(ctx)(all_i, all_d)
        torch.cuda.synchronize()
        assert np.array_equal(got_i.cpu().numpy(), want_i) and np.array_equal(got_d.cpu().numpy(), want_d)
        assert got_i.cpu().numpy()[0, 0] == 7 and got_i.cpu().numpy()[0, 1] == n // 2 + 3      # the tie, in id order
    finally:
        ctx.close()


def test_group_of_one_device_needs_no_nccl():
    n, dim, nq, k = 2000, 64, 100, 4
    db, q = make_db(n, dim, 91, nq)
    g = capi.Group(ndev=1)
    try:
        assert g.nranks == 1
        s = capi.NetvladDB(g.ctxs[0], dim, n)
        s.append(db)
        i, d = g.netvlad_topk([s], q, k, capi.NN_TENSOR)
        ref_i, ref_d = fm.topk(q, db, k)
        assert np.array_equal(i, ref_i) and np.array_equal(d, ref_d)
        assert g.last_gather_bytes == 0
    finally:
        g.close()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (NCCL all-gather between shards)")
@pytest.mark.parametrize("ndev", [2, 4, 8])
def test_group_sharded_search_over_nccl(ndev):
    if torch.cuda.device_count() < ndev:
        pytest.skip("fewer than %d GPUs" % ndev)
    assert capi.Group.nccl_version() > 0
    n, dim, nq, k = 20011, 128, 1500, 5
    db, q = make_db(n, dim, 92, nq)
    g = capi.Group(ndev=ndev)
    try:
        shards = []
        for r in range(ndev):
            lo, hi = sharding.shard_rows(n, r, ndev)
            s = capi.NetvladDB(g.ctxs[r], dim, hi - lo, base=lo)
            s.append(db[lo:hi])
            shards.append(s)
        ref_i, ref_d = fm.topk(q, db, k)
        for mode in (capi.NN_TENSOR, capi.NN_EXACT):
            i, d = g.netvlad_topk(shards, q, k, mode)
            assert np.array_equal(i, ref_i) and np.array_equal(d, ref_d)
        assert g.last_gather_bytes == nq * k * 12 * (ndev - 1) * ndev
    finally:
        g.close()
