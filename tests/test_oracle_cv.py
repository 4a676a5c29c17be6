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


# ---- round 2: P3P flags, four correspondences, n-point EPnP refit, refine loop (tests/golden/p3p.npz) ----------------
def _harness():
    import ctypes as C
    from tests import host_harness
    H = host_harness.load()
    return H, (lambda a: a.ctypes.data_as(C.c_void_p))


def test_p3p_solver_vs_cv2_golden(golden):
    """csrc/p3p.cuh (host build) against cv2.solvePnP(SOLVEPNP_P3P) on the golden 4-point sets.  Not an operation replay
    (DESIGN.md section 2): exact-to-rounding solvers agree to ~1e-9 on the pose; cv2's NaN answers count as no solution."""
    g = golden("p3p")
    H, dp = _harness()
    K = g["K"]
    K4 = np.array([K[0, 0], K[1, 1], K[0, 2], K[1, 2]])
    n_ok = 0
    for obj, img, ok, rv, tv in zip(g["p3p_obj"], g["p3p_img"], g["p3p_ok"], g["p3p_rvec"], g["p3p_tvec"]):
        R, t, r = np.zeros(9), np.zeros(3), np.zeros(3)
        got = H.hh_p3p4(dp(np.ascontiguousarray(obj)), dp(np.ascontiguousarray(img)), dp(K4), dp(R), dp(t), dp(r))
        assert bool(got) == bool(ok)
        if ok:
            assert np.abs(r - rv).max() < 1e-7 and np.abs(t - tv).max() < 1e-7
            n_ok += 1
    assert n_ok >= 55


def test_epnp_n_bit_exact_golden(golden):
    """The n-point EPnP of the P3P-flag refit (csrc/epnp.cuh:epnp_n, host build) is bit-identical to
    cv2.solvePnP(SOLVEPNP_EPNP) on f64 inputs, n = 4 ... 400."""
    g = golden("p3p")
    H, dp = _harness()
    K = g["K"]
    K4 = np.array([K[0, 0], K[1, 1], K[0, 2], K[1, 2]])
    for j in range(int(g["epnpn_n"])):
        obj, img = np.ascontiguousarray(g[f"epnpn_obj{j}"]), np.ascontiguousarray(g[f"epnpn_img{j}"])
        R, t, r = np.zeros(9), np.zeros(3), np.zeros(3)
        H.hh_epnp_n(dp(obj), dp(img), len(obj), dp(K4), dp(R), dp(t), dp(r))
        assert np.array_equal(r, g[f"epnpn_rvec{j}"]) and np.array_equal(t, g[f"epnpn_tvec{j}"]), j
        # and the pure-Python restatement (f32 inputs path differs: compare through cv2 on the same f32 data)
    obj32 = g["epnpn_obj3"].astype(np.float32)
    img32 = g["epnpn_img3"].astype(np.float32)
    ok, rv, tv = cv2.solvePnP(obj32, img32, K, None, flags=cv2.SOLVEPNP_EPNP)
    r2, t2, _ = cv_epnp.solve_epnp(obj32, img32, K[0, 0], K[1, 1], K[0, 2], K[1, 2])
    assert np.array_equal(np.array(r2), rv.ravel()) and np.array_equal(np.array(t2), tv.ravel())


def test_p3p_flag_ransac_restatement_matches_library(golden):
    g = golden("p3p")
    K = g["K"]
    z = np.zeros(3)
    for j in range(int(g["rs_n"])):
        obj, img = g[f"rs_obj{j}"], g[f"rs_img{j}"]
        res = pnp.restated_solve_pnp_ransac(obj, img, K, z, z, iterations=200, flags=2)
        assert res["ok"] == bool(g[f"rs_ok{j}"])
        assert np.array_equal(res["inliers"], g[f"rs_inl{j}"])
        if res["ok"]:
            assert np.abs(res["rvec"] - g[f"rs_rvec{j}"]).max() < 1e-9 and np.abs(res["tvec"] - g[f"rs_tvec{j}"]).max() < 1e-9
        # the refine loop on top of the default-flag RANSAC is deterministic (fixture written by the same function)
        ok, rv, tv, inl = pnp.cv_solve_pnp_ransac(obj, img, K, z, z, iterations=200)
        if ok and len(inl) >= 5:
            rv, tv, inl = pnp.refine_model(obj, img, K, rv, tv, inl, 5, 5)
        assert np.array_equal(inl, g[f"rf_inl{j}"]) and np.abs(rv - g[f"rf_rvec{j}"]).max() < 1e-9


def test_gao_p3p_restatement_solves_the_three_points():
    """oracle/cv_p3p.py (the P3P of OpenCV 3.2, Gao's quartic): every candidate it returns satisfies its three defining
    correspondences on well-conditioned sets, and the 4-point pick agrees with cv2 4.13 on most of them (the two solvers
    differ in the ill-conditioned tail, which is why K3 does not replay this one)."""
    from oracle import cv_p3p
    rng = np.random.default_rng(7)
    fx, fy, cx, cy = 500.0, 510.0, 320.0, 240.0
    K = pnp.K_matrix(fx, fy, cx, cy)
    close = 0
    for _ in range(200):
        obj = rng.uniform(-2, 2, (4, 3)).astype(np.float32)
        obj[:, 2] += 6
        rv, tv = rng.normal(0, 0.3, 3), rng.normal(0, 0.5, 3)
        img, _ = cv2.projectPoints(obj.astype(np.float64), rv, tv, K, None)
        img = img.reshape(-1, 2).astype(np.float32)
        ok, R, t = cv_p3p.solve_p3p(obj, img, fx, fy, cx, cy)
        ok2, rv2, tv2 = cv2.solvePnP(obj, img, K, None, flags=cv2.SOLVEPNP_P3P)
        if ok and ok2 and np.all(np.isfinite(tv2)):
            R2, _ = cv2.Rodrigues(rv2)
            close += int(np.abs(np.array(R) - R2).max() < 1e-4 and np.abs(np.array(t) - tv2.ravel()).max() < 1e-4)
    assert close >= 170
