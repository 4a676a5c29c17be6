"""K2 parity (GPU, through the C ABI): Hamming 2-NN, NNDR, uniqueness, correspondence list --
bit-exact against the oracle (cv2.BFMatcher) and the committed golden vectors."""
import numpy as np
import pytest

from multi_robot_slam_separators_b200 import capi, synth
from oracle import matching

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", params=["popc", "tensor"])
def ctx(request):
    """every test runs once per matcher kernel: hamming_knn.cu (POPC pipe) and hamming_mma.cu (tcgen05)"""
    c = capi.Context(0)
    c.set_match_kernel(request.param)
    yield c
    c.close()


def _store(ctx, descs, kp3d=None, max_feat=None):
    """descs: list of [n_i,32] arrays -> Frames with ragged counts."""
    F = max_feat or max(max(len(d) for d in descs), 1)
    n = len(descs)
    D = np.zeros((n, F, 32), np.uint8)
    K2 = np.zeros((n, F, 2), np.float32)
    K3 = np.ones((n, F, 3), np.float32)
    cnt = np.zeros(n, np.int32)
    for i, d in enumerate(descs):
        D[i, :len(d)] = d
        cnt[i] = len(d)
        if kp3d is not None:
            K3[i, :len(d)] = kp3d[i]
    f = capi.Frames(ctx, n, F)
    f.upload(0, D, K2, K3, cnt)
    return f


def _check_pair(out, p, q, t, kp3d_from, nndr, strict):
    idx, dist = matching.knn2(q, t)
    nq = len(q)
    assert np.array_equal(out["knn_idx"][p, :nq], idx)
    assert np.array_equal(out["knn_dist"][p, :nq], dist)
    pairs = matching.global_match(q, t, nndr, strict=strict)
    _, _, fi, ti = matching.pnp_correspondences(pairs, kp3d_from, np.zeros((max(nq, 1), 2), np.float32))
    m = out["n_matches"][p]
    assert m == len(fi)
    assert np.array_equal(out["match_from"][p, :m], fi.astype(np.uint16))
    assert np.array_equal(out["match_to"][p, :m], ti.astype(np.uint16))


def test_golden_vectors(ctx, golden):
    g = golden("hamming")
    fr = _store(ctx, [g["t"], g["ka_t"]], max_feat=128)
    to = _store(ctx, [g["q"], g["ka_q"]], max_feat=128)
    out = capi.match_pairs(ctx, fr, to, [[0, 0], [1, 1]], nndr=0.8)
    assert np.array_equal(out["knn_idx"][0, :96], g["idx"])
    assert np.array_equal(out["knn_dist"][0, :96], g["dist"])
    m = out["n_matches"][0]
    got = np.stack([out["match_from"][0, :m], out["match_to"][0, :m]], 1).astype(np.int32)
    assert np.array_equal(got, g["pairs"])
    # ties resolve to ascending trainIdx
    assert out["knn_idx"][1, 0].tolist() == [1, 3] and out["knn_dist"][1, 0].tolist() == [1, 1]
    out = capi.match_pairs(ctx, fr, to, [[0, 0]], nndr=0.8, strict=True)
    m = out["n_matches"][0]
    got = np.stack([out["match_from"][0, :m], out["match_to"][0, :m]], 1).astype(np.int32)
    assert np.array_equal(got, g["pairs_strict"])


def test_ragged_and_empty(ctx):
    rng = np.random.default_rng(7)
    sizes_from = [0, 1, 2, 5, 300, 1000, 1025, 33]
    sizes_to = [0, 1, 3, 1000, 257, 1024, 2, 1500]
    descs_f = [rng.integers(0, 256, (n, 32), dtype=np.uint8) for n in sizes_from]
    descs_t = [rng.integers(0, 256, (n, 32), dtype=np.uint8) for n in sizes_to]
    # make true matches so that NNDR passes for some, plus duplicates to exercise uniqueness
    for f, t in zip(descs_f, descs_t):
        k = min(len(f), len(t)) // 2
        t[:k] = f[:k]
        flips = rng.integers(0, 256, (k, 32), dtype=np.uint8) & rng.integers(0, 256, (k, 32), dtype=np.uint8) & \
            rng.integers(0, 256, (k, 32), dtype=np.uint8) & rng.integers(0, 256, (k, 32), dtype=np.uint8)
        t[:k] ^= flips
        if k > 4:
            t[k - 1] = t[0]  # two To rows on the same From row -> both dropped
    kp3 = [rng.standard_normal((n, 3)).astype(np.float32) for n in sizes_from]
    for k3 in kp3:
        if len(k3) > 3:
            k3[1, 2] = np.nan
            k3[3, 0] = np.inf
    fr = _store(ctx, descs_f, kp3, max_feat=1536)
    to = _store(ctx, descs_t, max_feat=1536)
    pairs = [[i, j] for i in range(len(sizes_from)) for j in range(len(sizes_to))]
    for strict in (False, True):
        out = capi.match_pairs(ctx, fr, to, pairs, nndr=0.8, strict=strict)
        for p, (i, j) in enumerate(pairs):
            _check_pair(out, p, descs_t[j], descs_f[i], kp3[i], 0.8, strict)


def test_synthetic_keyframes_config1_shape(ctx):
    A, B = synth.make_robots(2, 24, n_feat=1000, dim=128, seed=1001)
    fa, fb = capi.frames_from_robot(ctx, A), capi.frames_from_robot(ctx, B)
    rng = np.random.default_rng(3)
    same = [(a, b) for a in range(24) for b in range(24) if A.place[a] == B.place[b]]
    pairs = same[:8] + [(int(rng.integers(24)), int(rng.integers(24))) for _ in range(16)]
    out = capi.match_pairs(ctx, fa, fb, pairs, nndr=0.8)
    for p, (a, b) in enumerate(pairs):
        _check_pair(out, p, B.desc[b], A.desc[a], A.kp3d[a], 0.8, False)
    assert out["n_matches"][:len(same[:8])].min() > 100


def test_invalid_arguments(ctx):
    fr = _store(ctx, [np.zeros((4, 32), np.uint8)])
    with pytest.raises(capi.SepbError):
        capi.match_pairs(ctx, fr, fr, [[0, 5]])
    with pytest.raises(capi.SepbError):
        capi.Frames(ctx, 4, 5000)


def test_tensor_kernel_equals_popc_kernel_full_size():
    """size-independent property at BASELINE configs[1] frame size: both kernels agree bit for bit on
    every output (2-NN indices, distances, correspondence lists) over a few hundred 1000x1000 pairs."""
    A, B = synth.make_robots(2, 48, n_feat=1000, dim=128, seed=2002)
    rng = np.random.default_rng(11)
    A.count[5], A.count[9], B.count[7], B.count[11] = 0, 333, 0, 129      # ragged and empty frames inside the batch
    pairs = rng.integers(0, 48, (700, 2)).astype(np.int32)             # > 4 pairs per CTA of the persistent kernel
    outs = []
    # the tensor kernel in both accumulator-tile geometries (the launch takes the 256-column one for these frames;
    # SEPB_OPT_MATCH_TILE = 128 forces the geometry used for longer To frames) and both schedulers (persistent CTAs
    # walking their share of the pairs -- the default -- and one CTA per pair)
    for which, tn, sched in (("popc", 0, "persistent"), ("tensor", 0, "persistent"), ("tensor", 128, "persistent"),
                             ("tensor", 0, "per_pair"), ("tensor", 128, "per_pair")):
        c = capi.Context(0)
        c.set_match_kernel(which)
        c.set_match_tile(tn)
        c.set_match_scheduler(sched)
        fa, fb = capi.frames_from_robot(c, A), capi.frames_from_robot(c, B)
        outs.append(capi.match_pairs(c, fa, fb, pairs, nndr=0.8))
        fa.close(); fb.close(); c.close()
    for other in outs[1:]:
        for k in ("knn_idx", "knn_dist", "n_matches", "n_unique_to"):
            assert np.array_equal(outs[0][k], other[k]), k
        for p in range(len(pairs)):
            m = outs[0]["n_matches"][p]
            assert np.array_equal(outs[0]["match_from"][p, :m], other["match_from"][p, :m])
            assert np.array_equal(outs[0]["match_to"][p, :m], other["match_to"][p, :m])
