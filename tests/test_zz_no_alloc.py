"""sepb_netvlad_topk_exact_static_dev (include/sepb200.h): the exact NetVLAD search with its two-pass partial lists in a
file-scope device array instead of cudaMalloc'ed scratch -- no counterpart in the reference, added for callers that
cannot allocate (CUDA graph capture, serving loops).  Checked through the C ABI against the existing entry point
sepb_netvlad_topk (mode exact), bit for bit like tests/test_k1_netvlad_gpu.py, on calls in a row that share the array.
Named test_zz_* so that pytest collects it last."""
import numpy as np
import pytest

from multi_robot_slam_separators_b200 import capi

pytestmark = pytest.mark.gpu


def device_call(torch, db, q, k, row_mask=None):
    d_q = torch.from_numpy(np.ascontiguousarray(q, np.float64)).cuda()
    nq = q.shape[0]
    d_idx = torch.full((nq, k), -7, dtype=torch.int32, device="cuda")
    d_dist = torch.full((nq, k), -7.0, dtype=torch.float64, device="cuda")
    d_rm = torch.from_numpy(row_mask).cuda() if row_mask is not None else None
    torch.cuda.synchronize()          # torch's stream is not the context's
    db.topk_exact_static_dev(d_q.data_ptr(), nq, k, d_idx.data_ptr(), d_dist.data_ptr(),
                             d_rm.data_ptr() if d_rm is not None else 0)
    db.ctx.sync()
    return d_idx.cpu().numpy(), d_dist.cpu().numpy()


def test_static_scratch_search_matches_the_allocating_entry_point():
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(77)
    D, N = 128, 3000
    rows = rng.normal(size=(N, D))
    rows[100] = rows[7]                                   # exact ties -> lowest index first
    ctx = capi.Context(0)
    db = capi.NetvladDB(ctx, D, N)
    try:
        db.append(rows)
        db.set_col_mask([3, 500, 2999])
        db.topk_exact_static_dev(0, 0, 5, 0, 0)           # set-up only call (nq = 0)
        # call 1: 300 queries -> 3 row blocks x 64 database splits: 1.1 MB of partial lists in the static array
        q1 = rows[rng.integers(0, N, 300)] + 0.05 * rng.normal(size=(300, D))
        q1[0] = rows[7]
        rm1 = np.zeros(300, np.uint8)
        rm1[[4, 250]] = 1
        got_i, got_d = device_call(torch, db, q1, 5, rm1)
        want_i, want_d = db.topk(q1, k=5, mode=capi.NN_EXACT, local_used=np.nonzero(rm1)[0])
        assert np.array_equal(got_i, want_i) and np.array_equal(got_d, want_d)
        assert got_i[0, 0] == 7 and got_i[0, 1] == 100 and got_d[0, 0] == 0.0 and (got_i[4] == -1).all()
        # call 2 right behind it, other shape (fewer, wider lists: stale contents of call 1 lie under and beside them)
        q2 = rng.normal(size=(77, D))
        got_i2, got_d2 = device_call(torch, db, q2, 16)
        want_i2, want_d2 = db.topk(q2, k=16, mode=capi.NN_EXACT)
        assert np.array_equal(got_i2, want_i2) and np.array_equal(got_d2, want_d2)
        # call 3 = call 1 again: same answer after the array was overwritten
        got_i3, got_d3 = device_call(torch, db, q1, 5, rm1)
        assert np.array_equal(got_i3, want_i) and np.array_equal(got_d3, want_d)
        # a single split writes straight to the outputs (no scratch): many queries against few rows
        small = capi.NetvladDB(ctx, D, 40)
        small.append(rows[:20])
        got_i4, got_d4 = device_call(torch, small, q1, 3)
        want_i4, want_d4 = small.topk(q1, k=3, mode=capi.NN_EXACT)
        assert np.array_equal(got_i4, want_i4) and np.array_equal(got_d4, want_d4)
        small.close()
        # beyond the array: refused with a message, nothing launched
        big_q = torch.zeros((30000, D), dtype=torch.float64, device="cuda")
        out_i = torch.zeros((30000, 16), dtype=torch.int32, device="cuda")
        out_d = torch.zeros((30000, 16), dtype=torch.float64, device="cuda")
        with pytest.raises(capi.SepbError, match="static scratch holds"):
            db.topk_exact_static_dev(big_q.data_ptr(), 30000, 16, out_i.data_ptr(), out_d.data_ptr())
        with pytest.raises(capi.SepbError):
            db.topk_exact_static_dev(big_q.data_ptr(), 10, 17, out_i.data_ptr(), out_d.data_ptr())      # k > 16
    finally:
        db.close()
        ctx.close()
