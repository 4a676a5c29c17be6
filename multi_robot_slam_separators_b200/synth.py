"""Seeded synthetic stereo keyframes of the shapes BASELINE.json names.

Host-side data generation only (numpy); nothing here is on the compute path.

World: P places, each with its own landmark cloud inside the viewing frustum of the
place's canonical camera.  A keyframe = place pose composed with a small jitter; features =
projected landmarks (pixel noise, stereo-depth noise, descriptor bit flips) topped up with
clutter to exactly F features.  Robots visit overlapping subsets of places.  NetVLAD
descriptors are 4096-d unit vectors (place vector + noise) truncated to D WITHOUT
re-normalising (R/scripts/data_handler.py:157-158).
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

# KITTI-like rectified stereo rig
FX = FY = 718.856
CX, CY = 607.1928, 185.2157
WIDTH, HEIGHT = 1241, 376
BASELINE = 0.54
# optical (x right, y down, z fwd) -> base (x fwd, y left, z up): p_base = R_LOCAL @ p_cam
R_LOCAL = np.array([[0.0, 0.0, 1.0], [-1.0, 0.0, 0.0], [0.0, -1.0, 0.0]])


@dataclass
class RobotFrames:
    """Feature store of one robot, fixed stride F (counts may be smaller: ragged)."""
    desc: np.ndarray        # [N, F, 32] u8
    kp2d: np.ndarray        # [N, F, 2] f32 pixel
    kp3d: np.ndarray        # [N, F, 3] f32 base frame (NaN = stereo failure)
    octave: np.ndarray      # [N, F] i32
    count: np.ndarray       # [N] i32 valid features per frame
    netvlad: np.ndarray     # [N, D] f64 (values exactly representable in f32)
    place: np.ndarray       # [N] i64 place id
    pose: np.ndarray        # [N, 4, 4] f64 world <- base
    landmark: np.ndarray = field(default=None)  # [N, F] i32 landmark id within place, -1 clutter


def _rot_zyx(yaw, pitch, roll):
    cy, sy = np.cos(yaw), np.sin(yaw)
    cp, sp = np.cos(pitch), np.sin(pitch)
    cr, sr = np.cos(roll), np.sin(roll)
    R = np.empty(yaw.shape + (3, 3))
    R[..., 0, 0] = cy * cp
    R[..., 0, 1] = cy * sp * sr - sy * cr
    R[..., 0, 2] = cy * sp * cr + sy * sr
    R[..., 1, 0] = sy * cp
    R[..., 1, 1] = sy * sp * sr + cy * cr
    R[..., 1, 2] = sy * sp * cr - cy * sr
    R[..., 2, 0] = -sp
    R[..., 2, 1] = cp * sr
    R[..., 2, 2] = cp * cr
    return R


def place_assignment(n_robots, n_frames, overlap, rng):
    """Robot 0 visits places 0..n-1; robot r>0 revisits `overlap` of robot 0's places (random
    positions in its own sequence) and otherwise sees fresh places."""
    places = [np.arange(n_frames, dtype=np.int64)]
    nxt = n_frames
    for _ in range(1, n_robots):
        p = np.arange(nxt, nxt + n_frames, dtype=np.int64)
        nxt += n_frames
        k = int(round(overlap * n_frames))
        where = rng.choice(n_frames, size=k, replace=False)
        p[where] = rng.choice(n_frames, size=k, replace=False)
        places.append(p)
    return places


def make_netvlad(place_ids, dim, seed, noise=0.25, full_dim=4096, chunk=2048):
    """Unit 4096-d place vector + noise/sqrt(4096), re-normalised, truncated to `dim`."""
    n = len(place_ids)
    out = np.empty((n, dim), dtype=np.float64)
    for s in range(0, n, chunk):
        ids = place_ids[s:s + chunk]
        g = np.empty((len(ids), full_dim), dtype=np.float32)
        for i, p in enumerate(ids):
            g[i] = np.random.default_rng([seed, 7, int(p)]).standard_normal(full_dim, dtype=np.float32)
        g /= np.linalg.norm(g, axis=1, keepdims=True)
        nz = np.random.default_rng([seed, 11, s]).standard_normal(g.shape, dtype=np.float32)
        v = g + (noise / np.sqrt(full_dim)) * nz
        v /= np.linalg.norm(v, axis=1, keepdims=True)
        out[s:s + chunk] = v[:, :dim].astype(np.float32).astype(np.float64)
    return out


def make_frames(place_ids, robot, seed, n_feat=1000, dim=128, n_landmarks=700, flip_p=0.06,
                pix_sigma=0.5, disp_sigma=0.3, nan_frac=0.05, true_frac=0.55, gross_frac=0.35, chunk=256,
                hard_negatives=0.0, vocabulary=2000):
    """Generate one robot's keyframes for the given place ids.

    hard_negatives > 0: that fraction of the clutter features takes its descriptor from a vocabulary of `vocabulary`
    words shared by every place and robot (with the same bit-flip noise as landmarks), like repetitive structure seen by
    a real ORB front-end: frames of DIFFERENT places then share a few dozen words, pass the ratio test on them and send
    all-outlier correspondence sets into RANSAC, which burns every iteration (the K3-bound regime).  0 = every clutter
    descriptor is uniform random, so unrelated frames produce ~no matches (the configs[1] mix of round 1)."""
    n = len(place_ids)
    F = n_feat
    desc = np.empty((n, F, 32), dtype=np.uint8)
    kp2d = np.empty((n, F, 2), dtype=np.float32)
    kp3d = np.empty((n, F, 3), dtype=np.float32)
    octave = np.empty((n, F), dtype=np.int32)
    lmk = np.full((n, F), -1, dtype=np.int32)
    pose = np.empty((n, 4, 4), dtype=np.float64)
    L = n_landmarks
    for s in range(0, n, chunk):
        ids = np.asarray(place_ids[s:s + chunk])
        B = len(ids)
        # per-place landmarks (deterministic in the place id): canonical camera frame
        lm_cam = np.empty((B, L, 3))
        lm_desc = np.empty((B, L, 32), dtype=np.uint8)
        place_T = np.empty((B, 4, 4))
        for i, p in enumerate(ids):
            r = np.random.default_rng([seed, 3, int(p)])
            z = r.uniform(4.0, 40.0, L)
            u = r.uniform(-150.0, WIDTH + 150.0, L)
            v = r.uniform(-60.0, HEIGHT + 60.0, L)
            lm_cam[i, :, 0] = (u - CX) / FX * z
            lm_cam[i, :, 1] = (v - CY) / FY * z
            lm_cam[i, :, 2] = z
            lm_desc[i] = r.integers(0, 256, (L, 32), dtype=np.uint8)
            yaw = r.uniform(-np.pi, np.pi)
            T = np.eye(4)
            T[:3, :3] = _rot_zyx(np.array(yaw), np.array(0.0), np.array(0.0))
            T[:3, 3] = [r.uniform(-500, 500), r.uniform(-500, 500), 0.0]
            place_T[i] = T
        r = np.random.default_rng([seed, 5, robot, s])
        # jittered keyframe pose relative to the place base frame
        J = np.tile(np.eye(4), (B, 1, 1))
        J[:, :3, :3] = _rot_zyx(r.uniform(-0.17, 0.17, B), r.uniform(-0.03, 0.03, B), r.uniform(-0.03, 0.03, B))
        J[:, :3, 3] = np.stack([r.uniform(-1, 1, B), r.uniform(-1, 1, B), r.uniform(-0.1, 0.1, B)], axis=1)
        pose[s:s + B] = place_T @ J
        # landmarks: canonical cam -> place base -> jittered base -> jittered cam
        lm_base = lm_cam @ R_LOCAL.T                                   # [B,L,3]
        Rj, tj = J[:, :3, :3], J[:, :3, 3]
        lm_jb = (lm_base - tj[:, None, :]) @ Rj                       # R^T (p - t)
        lm_jc = lm_jb @ R_LOCAL                                        # R_LOCAL^T p
        z = lm_jc[..., 2]
        u = lm_jc[..., 0] / z * FX + CX
        v = lm_jc[..., 1] / z * FY + CY
        vis = (z > 1.0) & (u >= 0) & (u < WIDTH) & (v >= 0) & (v < HEIGHT)
        n_true_max = min(int(true_frac * F), L)   # a place has L landmarks: larger frames get more clutter
        score = r.random((B, L))
        score[~vis] = 2.0
        order = np.argsort(score, axis=1)[:, :n_true_max]             # random visible subset first
        sel_vis = np.take_along_axis(vis, order, axis=1)              # [B,n_true_max]
        # clutter everywhere, then overwrite with true landmarks
        zc = r.uniform(3.0, 60.0, (B, F))
        uc = r.uniform(0, WIDTH, (B, F))
        vc = r.uniform(0, HEIGHT, (B, F))
        dsc = r.integers(0, 256, (B, F, 32), dtype=np.uint8)
        if hard_negatives > 0:     # own random stream: the default data does not change
            r2 = np.random.default_rng([seed, 17, robot, s])
            vocab = np.random.default_rng([seed, 13]).integers(0, 256, (vocabulary, 32), dtype=np.uint8)
            hn = r2.random((B, F)) < hard_negatives
            word = r2.integers(0, vocabulary, (B, F))
            wflip = np.packbits(r2.random((B, F, 256)) < flip_p, axis=2, bitorder="little")
            dsc = np.where(hn[..., None], vocab[word] ^ wflip, dsc)
        lid = np.full((B, F), -1, dtype=np.int32)
        bi = np.arange(B)[:, None]
        zt = np.take_along_axis(z, order, axis=1)
        ut = np.take_along_axis(u, order, axis=1)
        vt = np.take_along_axis(v, order, axis=1)
        dt = lm_desc[bi, order]                                        # [B,n_true_max,32]
        flips = r.random((B, n_true_max, 256)) < flip_p
        dt = dt ^ np.packbits(flips, axis=2, bitorder="little")
        m = sel_vis
        zc[:, :n_true_max] = np.where(m, zt, zc[:, :n_true_max])
        uc[:, :n_true_max] = np.where(m, ut, uc[:, :n_true_max])
        vc[:, :n_true_max] = np.where(m, vt, vc[:, :n_true_max])
        dsc[:, :n_true_max] = np.where(m[..., None], dt, dsc[:, :n_true_max])
        lid[:, :n_true_max] = np.where(m, order, -1)
        # measurement noise: pixels and stereo disparity
        un = uc + r.normal(0, pix_sigma, (B, F))
        vn = vc + r.normal(0, pix_sigma, (B, F))
        disp = FX * BASELINE / zc + r.normal(0, disp_sigma, (B, F))
        # gross stereo mismatches: descriptor matches but the 3D point is wrong (RANSAC outliers)
        gross = r.random((B, F)) < gross_frac
        disp = np.where(gross, r.uniform(1.0, 100.0, (B, F)), disp)
        zn = FX * BASELINE / np.maximum(disp, 0.05)
        pc = np.stack([(un - CX) / FX * zn, (vn - CY) / FY * zn, zn], axis=-1)
        pb = pc @ R_LOCAL.T
        bad = r.random((B, F)) < nan_frac
        pb[bad] = np.nan
        # shuffle feature order per frame
        perm = np.argsort(r.random((B, F)), axis=1)
        desc[s:s + B] = dsc[bi, perm]
        kp2d[s:s + B, :, 0] = np.take_along_axis(un, perm, axis=1)
        kp2d[s:s + B, :, 1] = np.take_along_axis(vn, perm, axis=1)
        kp3d[s:s + B] = pb[bi, perm]
        lmk[s:s + B] = np.take_along_axis(lid, perm, axis=1)
        oc = r.integers(0, 8, (B, F), dtype=np.int32)                 # clutter: any pyramid level
        lm_now = lmk[s:s + B]
        # a landmark is detected at the same pyramid level in every view (pass 2 only pairs equal octaves,
        # R/src/myRegistrationVis.cpp:723-734); derived from its id so the random stream is unchanged
        oc_lm = ((lm_now.astype(np.int64) * 2654435761) >> 7) % 8
        octave[s:s + B] = np.where(lm_now >= 0, oc_lm, oc).astype(np.int32)
    netvlad = make_netvlad(np.asarray(place_ids), dim, seed)
    return RobotFrames(desc=desc, kp2d=kp2d, kp3d=kp3d, octave=octave,
                       count=np.full(n, F, dtype=np.int32), netvlad=netvlad,
                       place=np.asarray(place_ids, dtype=np.int64), pose=pose, landmark=lmk)


def make_robots(n_robots=2, n_frames=200, n_feat=1000, dim=128, overlap=0.3, seed=1000, **kw):
    """BASELINE.json config shapes: e.g. (2, 200, 1000, 128) = configs[0], (2, 10000, 1000, 128) = configs[1]."""
    rng = np.random.default_rng([seed, 1])
    places = place_assignment(n_robots, n_frames, overlap, rng)
    return [make_frames(places[r], r, seed, n_feat=n_feat, dim=dim, **kw) for r in range(n_robots)]


def relative_pose_gt(frames_from: RobotFrames, i_from: int, frames_to: RobotFrames, i_to: int):
    """Ground-truth T (From-base <- To-base), what estimate_transformation should recover."""
    return np.linalg.inv(frames_from.pose[i_from]) @ frames_to.pose[i_to]
