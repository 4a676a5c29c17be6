"""K1 parity (GPU, through the C ABI): NetVLAD distances bit-identical to scipy.cdist, candidate
ids exact, masks and the find_matches selection loop (R/scripts/data_handler.py:166-209)."""
import numpy as np
import pytest

from multi_robot_slam_separators_b200 import capi, synth
from oracle import find_matches as fm

pytestmark = pytest.mark.gpu

MODES = [capi.NN_EXACT, capi.NN_TENSOR]


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


@pytest.mark.parametrize("mode", MODES)
def test_append_only_and_empty(ctx, mode):
    rng = np.random.default_rng(1)
    b = rng.standard_normal((50, 128))
    db = capi.NetvladDB(ctx, 128, 64)
    idx, dist = db.topk(b[:4], k=1, mode=mode)
    assert (idx == -1).all() and np.isinf(dist).all()        # empty database
    db.append(b[:20])
    idx, dist = db.topk(b[:10], k=1, mode=mode)              # converts the first 20 rows (tensor mode)
    assert idx.ravel().tolist() == list(range(10)) and (dist == 0).all()
    db.append(b[20:50])
    assert db.size == 50
    idx, dist = db.topk(b[:40], k=1, mode=mode)              # incremental conversion of the appended rows
    assert idx.ravel().tolist() == list(range(40)) and (dist == 0).all()
    with pytest.raises(capi.SepbError):
        db.append(rng.standard_normal((20, 128)))            # capacity exceeded -> error, not abort


@pytest.mark.parametrize("mode", MODES)
def test_config1_find_matches(ctx, mode):
    A, B = synth.make_robots(2, 200, n_feat=4, dim=128, seed=1000)
    db = _db(ctx, B.netvlad)
    i1, d1 = db.topk(A.netvlad, k=1, mode=mode)
    got = capi.select_matches(i1, d1, 0.13, 20)
    want = fm.find_matches(A.netvlad, B.netvlad, 0.13, 20)
    assert got == want and len(got) == 20


# ---- the tcgen05 filter itself -----------------------------------------------------------------
def _coeff(D):
    """filter_coeff() of netvlad_gemm.cu"""
    Dp = (D + 63) // 64 * 64
    return 1.5 * (3 * 2.0 ** -18 + (3 * Dp / 16) * 2.0 ** -23 + 3 * 2.0 ** -24)


@pytest.mark.parametrize("shape,scale", [((300, 1000, 128), 0.016), ((257, 700, 128), 3.0), ((140, 600, 4096), 0.0156),
                                         ((130, 300, 77), 1e-3), ((64, 520, 512), 0.05), ((100, 400, 40), 0.1)])
def test_filter_error_within_bound(ctx, shape, scale):
    """The raw GEMM values v = |d|^2 - 2 q.d (split-bf16, f32 TMEM accumulators) must sit well inside the
    bound E = C(D) (|q|^2 + max|d|^2) the completeness proof relies on; all-positive data is the worst case
    for accumulation error."""
    nq, n, D = shape
    rng = np.random.default_rng(D + n)
    for positive in (False, True):
        q = rng.standard_normal((nq, D)) * scale
        b = rng.standard_normal((n, D)) * scale
        if positive:
            q, b = np.abs(q), np.abs(b)
        db = _db(ctx, b)
        v, idx, dist, stats = db.filter_debug(q, k=5)
        exact = (b * b).sum(1)[None, :] - 2.0 * q @ b.T
        bound = _coeff(D) * ((q * q).sum(1)[:, None] + (b * b).sum(1).max())
        ratio = np.abs(v.astype(np.float64) - exact) / bound
        assert np.isfinite(v).all()
        assert ratio.max() < 0.5, f"filter error {ratio.max():.3f} of the bound"
        ridx, rdist = fm.topk(q, b, 5)
        assert np.array_equal(idx, ridx) and np.array_equal(dist, rdist)
        assert stats[2] == 0 and stats[1] == nq                 # no row needed the exact scan
        assert stats[0] < 3 * 5 * nq                            # the threshold keeps the re-rank short
        db.close()


def test_filter_fallback_rows(ctx):
    """Rows the filter cannot prove complete take the exact scan and still match cdist: a cluster of
    duplicates larger than the per-row list, non-finite and huge entries, tiny norms."""
    rng = np.random.default_rng(5)
    D, n = 128, 900
    b = rng.standard_normal((n, D)) * 0.02
    b[100:140] = b[100]                       # 40 identical rows > list capacity 12
    q = rng.standard_normal((64, D)) * 0.02
    q[0] = b[100] + 1e-9                      # its top-5 are all inside the duplicate cluster
    q[1, 3] = np.nan
    q[2, 5] = np.inf
    q[3] = 0.0
    db = _db(ctx, b)
    v, idx, dist, stats = db.filter_debug(q, k=5)
    ridx, rdist = fm.topk(q, b, 5)
    ridx[1], rdist[1] = -1, np.inf            # NaN distances are never candidates (same rule as the exact scan)
    eidx, edist = db.topk(q, k=5, mode=capi.NN_EXACT)
    assert np.array_equal(eidx, ridx) and np.array_equal(edist, rdist)
    assert np.array_equal(idx, ridx) and np.array_equal(dist, rdist)
    assert idx[0].tolist() == [100, 101, 102, 103, 104]
    assert stats[2] >= 1
    db.close()
    # a database row that overflows f32 poisons the bound: every row falls back, results unchanged
    b2 = b.copy()
    b2[7, 0] = 1e30
    b2[9, 1] = np.nan
    db = _db(ctx, b2)
    idx, dist = db.topk(q[4:], k=3, mode=capi.NN_TENSOR)
    ridx, rdist = fm.topk(q[4:], b2, 3)
    assert np.array_equal(idx, ridx) and np.array_equal(dist, rdist)
    db.close()
    # tiny norms
    db = _db(ctx, b * 1e-20)
    idx, dist = db.topk(q[4:] * 1e-20, k=3, mode=capi.NN_TENSOR)
    ridx, rdist = fm.topk(q[4:] * 1e-20, b * 1e-20, 3)
    assert np.array_equal(idx, ridx) and np.array_equal(dist, rdist)
    db.close()


def test_tensor_mode_large(ctx):
    """configs[1]-like slice: synthetic truncated NetVLAD, 4000 x 10000, k = 5 and k = 1."""
    places = synth.place_assignment(2, 10000, 0.3, np.random.default_rng([2000, 1]))
    nv_a, nv_b = synth.make_netvlad(places[0], 128, 2000), synth.make_netvlad(places[1], 128, 2000)
    db = _db(ctx, nv_a)
    q = nv_b[:4000]
    for k in (5, 1, 8):
        idx, dist = db.topk(q, k=k, mode=capi.NN_TENSOR)
        eidx, edist = db.topk(q, k=k, mode=capi.NN_EXACT)
        assert np.array_equal(idx, eidx) and np.array_equal(dist, edist)
    ridx, rdist = fm.topk(q[:500], nv_a, 5)
    idx, dist = db.topk(q[:500], k=5, mode=capi.NN_TENSOR)
    assert np.array_equal(idx, ridx) and np.array_equal(dist, rdist)
    v, idx, dist, stats = db.filter_debug(q[:1000], k=5, want_v=False)
    assert stats[2] <= 10, f"{stats[2]} of 1000 rows fell back to the exact scan"
    db.close()
