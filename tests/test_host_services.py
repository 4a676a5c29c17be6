"""C++ host layer (host/separator_services.cpp): the reference's service handlers over POD message mirrors.
CPU part: it builds, loads and answers in-band failure without a CUDA device (never aborts).
GPU part: estimate_transformation / find_matches_compute / receive_separators against the batch path and the oracle."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from host import build as host_build  # noqa: E402
from multi_robot_slam_separators_b200 import build as lib_build, synth  # noqa: E402

vp, i32, f64 = C.c_void_p, C.c_int, C.c_double


def load():
    lib_build.build()
    L = C.CDLL(host_build.build())
    L.sh_device_create.restype = vp
    L.sh_device_create.argtypes = [i32]
    L.sh_device_ok.argtypes = [vp]
    L.sh_device_destroy.argtypes = [vp]
    L.sh_geo_create.restype = vp
    L.sh_geo_create.argtypes = [vp, f64, f64, f64, f64, i32, i32, i32, i32]
    L.sh_geo_destroy.argtypes = [vp]
    L.sh_geo_configure.argtypes = [vp, i32, i32, i32, vp, vp]
    L.sh_geo_configure.restype = None
    L.sh_geo_estimate.argtypes = [vp, vp, i32, vp, vp, vp, vp, i32, vp, vp, vp, vp, vp, vp, vp]
    L.sh_pr_create.restype = vp
    L.sh_pr_create.argtypes = [vp, i32, f64, i32, i32]
    L.sh_pr_destroy.argtypes = [vp]
    L.sh_pr_add_keyframe.argtypes = [vp, vp, i32, vp, i32, vp, vp, vp, i32]
    L.sh_pr_find_matches_compute.argtypes = [vp, vp, i32, vp, vp, vp, i32, vp]
    L.sh_pr_receive_separators.argtypes = [vp, i32, vp, vp, vp]
    L.sh_pr_dropped_int16.argtypes = [vp]
    L.sh_pr_last_error.argtypes = [vp]
    L.sh_pr_last_error.restype = C.c_char_p
    return L


def p(a):
    return C.c_void_p(a.ctypes.data)


def estimate(L, geo, R_from, i, R_to, j):
    nF, nT = int(R_from.count[i]), int(R_to.count[j])
    pose, cov = np.zeros(7), np.zeros(36)
    inl, mat = C.c_int(), C.c_int()
    ok = L.sh_geo_estimate(geo, p(R_from.desc[i]), nF, p(R_from.kp2d[i]), p(R_from.octave[i]), p(R_from.kp3d[i]),
                           p(R_to.desc[j]), nT, p(R_to.kp2d[j]), p(R_to.octave[j]), p(R_to.kp3d[j]), p(pose), p(cov),
                           C.byref(inl), C.byref(mat))
    return ok, pose, cov.reshape(6, 6), inl.value, mat.value


def test_builds_and_fails_in_band_without_device():
    import torch
    L = load()
    if torch.cuda.is_available():
        return
    dev = L.sh_device_create(0)
    assert L.sh_device_ok(dev) == 0
    geo = L.sh_geo_create(dev, synth.FX, synth.FY, synth.CX, synth.CY, synth.WIDTH, synth.HEIGHT, 5, 1024)
    A, B = synth.make_robots(2, 2, n_feat=64, dim=8, seed=3)
    ok, pose, cov, _, _ = estimate(L, geo, A, 0, B, 0)
    assert ok == 0 and not pose.any() and np.array_equal(cov, np.eye(6))       # success = false, zero pose, I6
    pr = L.sh_pr_create(dev, 8, 0.13, 20, 64)
    out = np.zeros(20, np.int16)
    assert L.sh_pr_find_matches_compute(pr, p(A.netvlad), A.netvlad.size, p(out), p(out), p(out), 20, None) == 0
    L.sh_pr_destroy(pr); L.sh_geo_destroy(geo); L.sh_device_destroy(dev)


def test_ros_node_compiles():
    """host/ros_node.cpp needs roscpp + tf + the generated message headers, none of which this image has: it is compiled
    (parsed and type-checked, -fsyntax-only) against the stand-in declarations of tests/host_harness/ros_stub, whose
    message structs mirror R/msg/*.msg and R/srv/*.srv field for field.  Catches what an #ifdef'ed-out file hides:
    a name clash with the generated multi_robot_separators:: types, calls of handlers that do not exist."""
    import subprocess
    stub = os.path.join(ROOT, "tests", "host_harness", "ros_stub")
    r = subprocess.run(["g++", "-std=c++17", "-fsyntax-only", "-DSEPB_WITH_ROS", "-I", stub, "-I", os.path.join(ROOT, "host"),
                        "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "host", "ros_node.cpp")],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    src = open(os.path.join(ROOT, "host", "ros_node.cpp")).read()
    for needle in ("place.addKeyframe(", "add_separators_pose_graph", "lookupTransform(", "geo.setLocalTransform("):
        assert needle in src


@pytest.mark.gpu
def test_estimate_transformation_matches_batch_path():
    from multi_robot_slam_separators_b200 import capi
    L = load()
    N = 30
    A, B = synth.make_robots(2, N, n_feat=1000, dim=8, seed=321, overlap=0.5)
    pairs = [(o, l) for l in range(N) for o in range(N) if A.place[o] == B.place[l]][:6] + [(1, 0)]
    ctx = capi.Context(0)
    fa, fb = capi.frames_from_robot(ctx, A), capi.frames_from_robot(ctx, B)
    prm = capi.pnp_params(synth.FX, synth.FY, synth.CX, synth.CY)
    ref = capi.estimate_transformations(ctx, fa, fb, np.array(pairs, np.int32), prm,
                                        capi.guess_params(synth.WIDTH, synth.HEIGHT), 0.8)
    dev = L.sh_device_create(0)
    assert L.sh_device_ok(dev) == 1
    geo = L.sh_geo_create(dev, synth.FX, synth.FY, synth.CX, synth.CY, synth.WIDTH, synth.HEIGHT, 5, 1000)
    n_ok = 0
    for k, (o, l) in enumerate(pairs):
        ok, pose, cov, inl, mat = estimate(L, geo, A, o, B, l)
        assert ok == int(ref["status"][k] == 1)
        T = ref["transform"][k].reshape(3, 4).astype(np.float64)
        if not ok:
            assert not pose.any()                                         # zero pose (MsgConversion.cpp:77-80)
            continue
        n_ok += 1
        assert inl == ref["n_inliers"][k] and mat == ref["n_matches"][k]
        assert np.array_equal(pose[:3], T[:, 3])
        x, y, z, w = pose[3:]
        Rq = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                       [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                       [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
        assert w >= 0 and abs(x * x + y * y + z * z + w * w - 1) < 1e-6 and np.abs(Rq - T[:, :3]).max() < 1e-6
        assert np.array_equal(cov, ref["covariance"][k].reshape(6, 6))
    assert n_ok >= 5
    L.sh_geo_destroy(geo); L.sh_device_destroy(dev)
    fa.close(); fb.close(); ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("flags,refine,cv32", [(2, 0, 0), (0, 3, 0), (0, 0, 1)])
def test_estimate_transformation_solver_options(flags, refine, cv32):
    """The handler's Vis/PnPFlags / Vis/PnPRefineIterations / OpenCV-3.2 switches reach the kernels: same answers as the
    batch C ABI with the same sepb_pnp_params, and different from the default where the option matters."""
    from multi_robot_slam_separators_b200 import capi
    L = load()
    N = 24
    A, B = synth.make_robots(2, N, n_feat=1000, dim=8, seed=322, overlap=0.5)
    pairs = [(o, l) for l in range(N) for o in range(N) if A.place[o] == B.place[l]][:5]
    ctx = capi.Context(0)
    fa, fb = capi.frames_from_robot(ctx, A), capi.frames_from_robot(ctx, B)
    gp = capi.guess_params(synth.WIDTH, synth.HEIGHT)
    prm = capi.pnp_params(synth.FX, synth.FY, synth.CX, synth.CY, flags=flags, refine_iterations=refine, minimal_model=bool(cv32))
    ref = capi.estimate_transformations(ctx, fa, fb, np.array(pairs, np.int32), prm, gp, 0.8)
    base = capi.estimate_transformations(ctx, fa, fb, np.array(pairs, np.int32),
                                         capi.pnp_params(synth.FX, synth.FY, synth.CX, synth.CY), gp, 0.8)
    dev = L.sh_device_create(0)
    geo = L.sh_geo_create(dev, synth.FX, synth.FY, synth.CX, synth.CY, synth.WIDTH, synth.HEIGHT, 5, 1000)
    L.sh_geo_configure(geo, flags, refine, cv32, None, None)
    n_ok = n_diff = 0
    for k, (o, l) in enumerate(pairs):
        ok, pose, cov, inl, mat = estimate(L, geo, A, o, B, l)
        assert ok == int(ref["status"][k] == 1)
        if ok:
            n_ok += 1
            T = ref["transform"][k].reshape(3, 4).astype(np.float64)
            assert np.array_equal(pose[:3], T[:, 3]) and inl == ref["n_inliers"][k] and mat == ref["n_matches"][k]
            n_diff += int(not np.array_equal(ref["transform"][k], base["transform"][k]))
    assert n_ok >= 4 and n_diff >= 1
    L.sh_geo_destroy(geo); L.sh_device_destroy(dev)
    fa.close(); fb.close(); ctx.close()


@pytest.mark.gpu
def test_estimate_transformation_with_bundle_adjustment():
    """Vis/BundleAdjustment = 1 on the handler (GeometricTools::setBundleAdjustment): the service answers what the batch
    C ABI gives with sepb_ctx_set_bundle_adjustment -- the adjusted pose of the second call and its inlier count."""
    from multi_robot_slam_separators_b200 import capi
    L = load()
    L.sh_geo_bundle_adjustment.argtypes = [C.c_void_p, C.c_int32, C.c_double]
    L.sh_geo_bundle_adjustment.restype = C.c_int32
    N = 24
    A, B = synth.make_robots(2, N, n_feat=1000, dim=8, seed=322, overlap=0.5)
    pairs = [(o, l) for l in range(N) for o in range(N) if A.place[o] == B.place[l]][:5]
    ctx = capi.Context(0)
    fa, fb = capi.frames_from_robot(ctx, A), capi.frames_from_robot(ctx, B)
    gp = capi.guess_params(synth.WIDTH, synth.HEIGHT)
    prm = capi.pnp_params(synth.FX, synth.FY, synth.CX, synth.CY)
    base = capi.estimate_transformations(ctx, fa, fb, np.array(pairs, np.int32), prm, gp, 0.8)
    ctx.set_bundle_adjustment(capi.ba_params(synth.BASELINE * synth.FY / synth.FX))
    ref = capi.estimate_transformations(ctx, fa, fb, np.array(pairs, np.int32), prm, gp, 0.8)
    dev = L.sh_device_create(0)
    geo = L.sh_geo_create(dev, synth.FX, synth.FY, synth.CX, synth.CY, synth.WIDTH, synth.HEIGHT, 5, 1000)
    assert L.sh_geo_bundle_adjustment(geo, 2, synth.BASELINE) == 0        # cvsba: not built, refused
    assert L.sh_geo_bundle_adjustment(geo, 1, synth.BASELINE) == 1
    n_ok = n_diff = 0
    for k, (o, l) in enumerate(pairs):
        ok, pose, cov, inl, mat = estimate(L, geo, A, o, B, l)
        assert ok == int(ref["status"][k] == 1)
        if ok:
            n_ok += 1
            T = ref["transform"][k].reshape(3, 4).astype(np.float64)
            assert np.array_equal(pose[:3], T[:, 3]) and inl == ref["n_inliers"][k] and mat == ref["n_matches"][k]
            n_diff += int(not np.array_equal(ref["transform"][k], base["transform"][k]))
    assert n_ok >= 4 and n_diff >= 1
    assert L.sh_geo_bundle_adjustment(geo, 0, 0.0) == 1
    L.sh_geo_destroy(geo); L.sh_device_destroy(dev)
    fa.close(); fb.close(); ctx.close()


@pytest.mark.gpu
def test_find_matches_compute_and_receive_separators():
    from oracle import find_matches as fm
    L = load()
    N = 200
    places = synth.place_assignment(2, N, 0.3, np.random.default_rng([1000, 1]))
    nv_a, nv_b = synth.make_netvlad(places[0], 128, 1000), synth.make_netvlad(places[1], 128, 1000)
    feats = synth.make_frames(places[1][:1], 1, 1000, n_feat=32, dim=8)          # the same small feature set for every frame
    dev = L.sh_device_create(0)
    pr = L.sh_pr_create(dev, 128, 0.13, 20, 1024)
    for i in range(N):                                                           # computing robot B's kept keyframes
        L.sh_pr_add_keyframe(pr, p(nv_b[i]), 128, p(feats.desc[0]), 32, p(feats.kp2d[0]), p(feats.octave[0]),
                             p(feats.kp3d[0]), 1000 + i)
    kf, lo, ot = np.zeros(64, np.int16), np.zeros(64, np.int16), np.zeros(64, np.int16)
    rows = C.c_int()
    # the querying robot sends its descriptors in two batches (append-only database)
    n1 = L.sh_pr_find_matches_compute(pr, p(nv_a[:120]), 120 * 128, p(kf), p(lo), p(ot), 64, C.byref(rows))
    want = fm.find_matches(nv_b, nv_a[:120], 0.13, 20)
    assert [(int(a), int(b)) for a, b in zip(lo[:n1], ot[:n1])] == want and rows.value == 32
    n2 = L.sh_pr_find_matches_compute(pr, p(nv_a[120:]), 80 * 128, p(kf), p(lo), p(ot), 64, C.byref(rows))
    want = fm.find_matches(nv_b, nv_a, 0.13, 20)
    got = [(int(a), int(b)) for a, b in zip(lo[:n2], ot[:n2])]
    assert got == want and len(got) >= 10
    assert kf[:n2].tolist() == [1000 + a for a, _ in got]                        # kf ids of the computing robot
    # separators come back: the first three succeeded, the next two failed (data_handler.py:373-419)
    fr = np.array([b for _, b in got[:5]], np.int16)
    to = np.array([a for a, _ in got[:5]], np.int16)
    succ = np.array([1, 1, 1, 0, 0], np.uint8)
    assert L.sh_pr_receive_separators(pr, 5, p(fr), p(to), p(succ)) == 1
    n3 = L.sh_pr_find_matches_compute(pr, p(nv_a[:0]), 0, p(kf), p(lo), p(ot), 64, None)
    want = fm.find_matches(nv_b, nv_a, 0.13, 20, local_used=to[:3].tolist(), other_used=fr[:3].tolist(),
                           pairs_ignored=[(int(to[3]), int(fr[3])), (int(to[4]), int(fr[4]))])
    assert [(int(a), int(b)) for a, b in zip(lo[:n3], ot[:n3])] == want
    L.sh_pr_destroy(pr); L.sh_device_destroy(dev)


@pytest.mark.gpu
def test_find_matches_compute_int16_wire_cap():
    """FindMatches.srv:3-5 carries int16 ids: a match whose database row does not fit is left out of the response
    (in-band, like "no match") and reported, never wrapped to a negative id."""
    L = load()
    rng = np.random.default_rng(5)
    n_db, D = 33000, 16
    db = rng.normal(size=(n_db, D)) * 0.3
    local = np.stack([db[32900] + 1e-3, db[100] + 1e-3, db[32767] + 1e-3])       # nearest rows: 32900 (too big), 100, 32767 (last that fits)
    feats = synth.make_frames(np.zeros(1, np.int64), 1, 1000, n_feat=32, dim=8)
    dev = L.sh_device_create(0)
    pr = L.sh_pr_create(dev, D, 0.13, 20, n_db)
    for i in range(3):
        L.sh_pr_add_keyframe(pr, p(np.ascontiguousarray(local[i])), D, p(feats.desc[0]), 32, p(feats.kp2d[0]), p(feats.octave[0]),
                             p(feats.kp3d[0]), 7 + i)
    kf, lo, ot = np.zeros(8, np.int16), np.zeros(8, np.int16), np.zeros(8, np.int16)
    n = L.sh_pr_find_matches_compute(pr, p(np.ascontiguousarray(db)), db.size, p(kf), p(lo), p(ot), 8, None)
    got = sorted((int(a), int(b)) for a, b in zip(lo[:n], ot[:n]))
    assert got == [(1, 100), (2, 32767)]
    assert L.sh_pr_dropped_int16(pr) == 1 and b"int16" in L.sh_pr_last_error(pr)
    L.sh_pr_destroy(pr); L.sh_device_destroy(dev)
