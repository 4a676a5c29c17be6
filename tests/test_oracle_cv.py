"""Oracle restatements vs the committed golden vectors (produced by cv2/scipy, see
tests/golden/make_golden.py) -- CPU only."""
import numpy as np
import pytest

from oracle import cvjacobi, cv_epnp, matching, pnp, find_matches as fm

cv2 = pytest.importorskip("cv2")


def test_jacobi_svd_bit_exact(golden):
    g = golden("svd")
    for i in range(int(g["n"])):
        w, u, vt = cvjacobi.svd(g[f"A{i}"])
        assert np.array_equal(w, g[f"w{i}"]) and np.array_equal(u, g[f"u{i}"]) and np.array_equal(vt, g[f"vt{i}"])


def test_cv_rng_known_answer():
    # first raw outputs of cv::RNG(0xFFFFFFFFFFFFFFFF) and first 5-point sets for n=357
    r = cvjacobi.CvRNG()
    assert [r.next() for _ in range(6)] == [130063605, 3133359004, 2578348940, 925327173, 1080261831, 2946015512]
    s = pnp.sample_index_stream(357, 5, 4)
    assert s.tolist() == [[294, 349, 335, 309, 180], [104, 288, 41, 188, 114], [214, 274, 99, 118, 8],
                          [126, 290, 260, 101, 69]]


def test_epnp_bit_exact(golden):
    g = golden("epnp")
    K = g["K"]
    for i in range(len(g["obj"])):
        rv, tv, _ = cv_epnp.solve_epnp(g["obj"][i], g["img"][i], K[0, 0], K[1, 1], K[0, 2], K[1, 2])
        assert np.array_equal(np.array(rv), g["rvec"][i])
        assert np.array_equal(np.array(tv), g["tvec"][i])


def test_hamming_knn2_golden(golden):
    g = golden("hamming")
    for knn in (matching.knn2, matching.knn2_numpy):
        idx, dist = knn(g["q"], g["t"])
        assert np.array_equal(idx, g["idx"]) and np.array_equal(dist, g["dist"])
    # ties -> ascending trainIdx
    assert g["ka"].tolist() == [[1, 1], [3, 1], [5, 1]]
    idx, dist = matching.knn2_numpy(g["ka_q"], g["ka_t"])
    assert idx[0].tolist() == [1, 3] and dist[0].tolist() == [1, 1]
    assert np.array_equal(np.array(matching.global_match(g["q"], g["t"], 0.8, knn=matching.knn2_numpy), np.int32).reshape(-1, 2), g["pairs"])
    assert np.array_equal(np.array(matching.global_match(g["q"], g["t"], 0.8, strict=True), np.int32).reshape(-1, 2), g["pairs_strict"])


def test_hamming_edge_cases():
    e = np.zeros((0, 32), np.uint8)
    one = np.zeros((1, 32), np.uint8)
    q = np.ones((3, 32), np.uint8)
    assert matching.knn2(e, q)[0].shape == (0, 2)
    assert (matching.knn2(q, e)[0] == -1).all()
    idx, dist = matching.knn2(q, one)            # k > #train: one neighbour only
    assert (idx[:, 0] == 0).all() and (idx[:, 1] == -1).all()
    assert matching.global_match(q, one) == []   # < 2 neighbours => rejected


def test_find_matches_golden(golden):
    g = golden("netvlad")
    d = fm.distance_matrix(g["local"], g["recv"])
    assert np.array_equal(d, g["d"])
    m = fm.find_matches(g["local"], g["recv"], 0.13, 20)
    assert len(m) > 3 and len({b for _, b in m}) == len(m)
    m2 = fm.find_matches(g["local"], g["recv"], 0.13, 20, g["local_used"], g["other_used"],
                         [tuple(p) for p in g["pairs_ignored"]])
    assert all(a not in g["local_used"] and b not in g["other_used"] for a, b in m2)
    assert all((a, b) not in {tuple(p) for p in g["pairs_ignored"]} for a, b in m2)
    idx, dist = fm.topk(g["local"], g["recv"], 5)
    assert np.array_equal(idx[:, 0], np.argmin(g["d"], axis=1))
    assert np.array_equal(dist, np.sort(g["d"], axis=1)[:, :5])
    # slot-consumption quirk (data_handler.py:194-200) and break-at-threshold
    assert fm.select_matches(np.array([0, 0, 1]), np.array([0.01, 0.02, 0.03]), 0.13, 2) == [(0, 0)]
    assert fm.select_matches(np.array([0, 1, 2]), np.array([0.01, 0.2, 0.03]), 0.13, 3) == [(0, 0), (2, 2)]
    assert fm.select_matches(np.array([0, 1]), np.array([0.13, 0.01]), 0.13, 5) == [(1, 1)]


def test_pnp_ransac_restatement_matches_library(golden):
    g = golden("pnp_ransac")
    K, rv0, tv0 = g["K"], g["rv0"], g["tv0"]
    for i in range(int(g["n"])):
        obj, img, iters = g[f"obj{i}"], g[f"img{i}"], int(g[f"iters{i}"])
        pairs = matching.global_match(g[f"desc_to{i}"], g[f"desc_from{i}"], 0.8)
        assert np.array_equal(np.array(pairs, np.int32), g[f"pairs{i}"])
        o2, i2, _, _ = matching.pnp_correspondences(pairs, g[f"kp3d_from{i}"], g[f"kp2d_to{i}"])
        assert np.array_equal(o2, obj) and np.array_equal(i2, img)
        res = pnp.restated_solve_pnp_ransac(obj, img, K, rv0, tv0, iterations=iters,
                                            sample_indices=g[f"samples{i}"], refit="py")
        assert res["ok"] == bool(g[f"ok{i}"])
        assert np.array_equal(res["inliers"], g[f"inliers{i}"])
        assert np.abs(res["rvec"] - g[f"rvec{i}"]).max() < 1e-9
        assert np.abs(res["tvec"] - g[f"tvec{i}"]).max() < 1e-9


def test_pnp_pure_python_hypotheses(golden):
    """One golden case with the pure-Python EPnP as the hypothesis solver (slow path)."""
    g = golden("pnp_ransac")
    i = 1
    res = pnp.restated_solve_pnp_ransac(g[f"obj{i}"], g[f"img{i}"], g["K"], g["rv0"], g["tv0"],
                                        iterations=int(g[f"iters{i}"]), hyp_solver="py")
    assert np.array_equal(res["inliers"], g[f"inliers{i}"])
