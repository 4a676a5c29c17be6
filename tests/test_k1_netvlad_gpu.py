"""K1 parity (GPU, through the C ABI): NetVLAD distances bit-identical to scipy.cdist, candidate
ids exact, masks and the find_matches selection loop (R/scripts/data_handler.py:166-209)."""
import numpy as np
import pytest

from multi_robot_slam_separators_b200 import capi, synth
from oracle import find_matches as fm

pytestmark = pytest.mark.gpu

MODES = [capi.NN_EXACT]


@pytest.fixture(scope="module")
def ctx():
    c = capi.Context(0)
    yield c
    c.close()


def _db(ctx, rows, base=0):
    db = capi.NetvladDB(ctx, rows.shape[1], max(len(rows), 1), base)
    if len(rows):
        db.append(rows)
    return db


@pytest.mark.parametrize("mode", MODES)
def test_golden_and_masks(ctx, golden, mode):
    g = golden("netvlad")
    local, recv = g["local"], g["recv"]
    db = _db(ctx, recv)
    idx, dist = db.topk(local, k=5, mode=mode)
    ridx, rdist = fm.topk(local, recv, 5)
    assert np.array_equal(idx, ridx) and np.array_equal(dist, rdist)      # bit-identical f64
    assert np.array_equal(dist, np.sort(g["d"], axis=1)[:, :5])
    # reference selection (data_handler.py:191-205)
    i1, d1 = db.topk(local, k=1, mode=mode)
    assert capi.select_matches(i1, d1, 0.13, 20) == fm.find_matches(local, recv, 0.13, 20)
    # masks: rows, columns, cells
    lu, ou, pi = g["local_used"], g["other_used"], [tuple(p) for p in g["pairs_ignored"]]
    db.set_col_mask(ou)
    idx, dist = db.topk(local, k=5, mode=mode, local_used=lu, pairs_ignored=pi)
    ridx, rdist = fm.topk(local, recv, 5, lu, ou, pi)
    assert np.array_equal(idx, ridx) and np.array_equal(dist, rdist)
    i1, d1 = db.topk(local, k=1, mode=mode, local_used=lu, pairs_ignored=pi)
    assert capi.select_matches(i1, d1, 0.13, 20) == fm.find_matches(local, recv, 0.13, 20, lu, ou, pi)
    db.set_col_mask(None)


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("shape", [(200, 200, 128), (1, 1, 128), (3, 700, 128), (300, 45, 77), (130, 513, 4096),
                                   (2000, 3000, 128)])
def test_shapes_vs_cdist(ctx, mode, shape):
    nq, n, D = shape
    rng = np.random.default_rng(nq * 7 + n)
    q = (rng.standard_normal((nq, D)) * 0.02).astype(np.float32).astype(np.float64)
    b = (rng.standard_normal((n, D)) * 0.02).astype(np.float32).astype(np.float64)
    if n > 4:
        b[3] = b[1]            # exact ties -> lowest index first
        q[0] = b[1]
    db = _db(ctx, b, base=1000)
    k = min(5, n)
    idx, dist = db.topk(q, k=5, mode=mode)
    ridx, rdist = fm.topk(q, b, 5)
    assert np.array_equal(idx[:, :k], ridx[:, :k] + 1000)
    assert np.array_equal(dist[:, :k], rdist[:, :k])
    assert (idx[:, k:] == -1).all() and np.isinf(dist[:, k:]).all()


def test_append_only_and_empty(ctx):
    rng = np.random.default_rng(1)
    b = rng.standard_normal((50, 128))
    db = capi.NetvladDB(ctx, 128, 64)
    idx, dist = db.topk(b[:4], k=1)
    assert (idx == -1).all() and np.isinf(dist).all()        # empty database
    db.append(b[:20]); db.append(b[20:50])
    assert db.size == 50
    idx, dist = db.topk(b[:10], k=1)
    assert idx.ravel().tolist() == list(range(10)) and (dist == 0).all()
    with pytest.raises(capi.SepbError):
        db.append(rng.standard_normal((20, 128)))            # capacity exceeded -> error, not abort


def test_config1_find_matches(ctx):
    A, B = synth.make_robots(2, 200, n_feat=4, dim=128, seed=1000)
    db = _db(ctx, B.netvlad)
    i1, d1 = db.topk(A.netvlad, k=1)
    got = capi.select_matches(i1, d1, 0.13, 20)
    want = fm.find_matches(A.netvlad, B.netvlad, 0.13, 20)
    assert got == want and len(got) == 20
