"""Seeded 2-view bundle adjustment problems shared by tests/test_oracle_ba.py, tests/test_ba_gpu.py and
tests/golden/make_golden.py: a stereo rig (KITTI-like intrinsics) sees N landmarks from two poses; pixels with Gaussian
noise, 3-D points from noisy disparities (base frames), a share of the To frame's 3-D points NaN (stereo failure ->
mono edges), a few gross pixel outliers among the 'inliers', and a perturbed PnP transform to start from."""
import numpy as np

FX, FY, CX, CY = 718.856, 718.856, 607.19, 185.22
BASELINE = 0.54
LOCAL_R = np.array([[0.0, 0.0, 1.0], [-1.0, 0.0, 0.0], [0.0, -1.0, 0.0]])   # optical -> base


def rodrigues(rv):
    th = float(np.linalg.norm(rv))
    if th < 1e-12:
        return np.eye(3)
    k = np.asarray(rv) / th
    Kx = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.eye(3) + np.sin(th) * Kx + (1 - np.cos(th)) * (Kx @ Kx)


def make(seed, n, n_out=4, nan_frac=0.1, local_t=(0.0, 0.0, 0.0), t_noise=0.03, r_noise=0.003, px_noise=0.5):
    rng = np.random.default_rng(seed)
    lt = np.asarray(local_t, np.float64)
    pc = np.stack([rng.uniform(-8, 8, n), rng.uniform(-2, 2, n), rng.uniform(4, 30, n)], 1)   # From camera frame
    pw = pc @ LOCAL_R.T + lt                                                                 # From base = world
    R = rodrigues(rng.normal(0, 0.05, 3))
    t = rng.normal(0, 0.5, 3)
    R2, t2 = R @ LOCAL_R, R @ lt + t                                                          # To camera in world
    pc2 = (pw - t2) @ R2

    def proj(p):
        return np.stack([FX * p[:, 0] / p[:, 2] + CX, FY * p[:, 1] / p[:, 2] + CY], 1)

    def noisy3(p):
        d = FX * BASELINE / p[:, 2] + rng.normal(0, 0.3, n)
        return p * ((FX * BASELINE / d) / p[:, 2])[:, None]

    k1 = (proj(pc) + rng.normal(0, px_noise, (n, 2))).astype(np.float32)
    k2 = (proj(pc2) + rng.normal(0, px_noise, (n, 2))).astype(np.float32)
    p3f = (noisy3(pc) @ LOCAL_R.T + lt).astype(np.float32)
    p3t = (noisy3(pc2) @ LOCAL_R.T + lt).astype(np.float32)          # To base frame
    p3t[rng.permutation(n)[: int(n * nan_frac)]] = np.nan
    out = rng.permutation(n)[:n_out]
    k2[out] += rng.choice([-1, 1], (n_out, 2)).astype(np.float32) * rng.uniform(15, 40, (n_out, 2)).astype(np.float32)
    T = np.concatenate([R, t[:, None]], 1)
    T0 = T.copy()
    T0[:, :3] = rodrigues(rng.normal(0, r_noise, 3)) @ R
    T0[:, 3] += rng.normal(0, t_noise, 3)
    cov = np.diag([rng.uniform(2e-4, 5e-3)] * 3 + [rng.uniform(1e-5, 1e-3)] * 3)
    return dict(T0=T0.astype(np.float32), T_true=T, kp2d_from=k1, kp3d_from=p3f, kp2d_to=k2, kp3d_to=p3t, cov=cov,
                local_t=lt, outliers=np.sort(out))


def default_set():
    """(seed, n, kwargs) of the committed fixture tests/golden/ba.npz."""
    return [(1, 40, {}), (2, 120, {}), (3, 300, dict(n_out=12)), (4, 25, dict(n_out=0, nan_frac=0.0)),
            (5, 80, dict(nan_frac=0.5)), (6, 60, dict(local_t=(0.2, -0.05, 0.4))), (7, 8, dict(n_out=2)),
            (8, 200, dict(t_noise=0.2, r_noise=0.02))]
