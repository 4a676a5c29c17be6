"""TEST INFRASTRUCTURE ONLY (oracle) -- never imported by the product path.

CPU restatement of the reference's global feature matching block
(R/src/myRegistrationVis.cpp:826-895) in its brute-force mode (Vis/CorNNType=3):
RTAB-Map `VWDictionary::addNewWords(descriptorsTo, 2)` (unvendored) is a
2-NN search of every To descriptor among the From descriptors with
`cv::BFMatcher(NORM_HAMMING).knnMatch(k=2)` -- executed here with cv2 4.13.0 -- an NNDR test
and then the uniqueness filter of :856-894.

Also restates how the surviving matches are ordered into the PnP correspondence list
(SURVEY.md A.2): ascending From index, rows with a non-finite From 3D point dropped.
"""
from __future__ import annotations

import numpy as np

try:  # cv2 is test/baseline infrastructure only
    import cv2
except Exception:  # pragma: no cover
    cv2 = None


def knn2(desc_to, desc_from):
    """For every To row: (idx1, d1, idx2, d2) best-2 among From rows.  Missing -> (-1, -1).

    cv2.BFMatcher ties resolve to the lowest trainIdx (pinned in tests)."""
    q = np.ascontiguousarray(desc_to, dtype=np.uint8)
    t = np.ascontiguousarray(desc_from, dtype=np.uint8)
    nq = q.shape[0]
    idx = np.full((nq, 2), -1, dtype=np.int32)
    dist = np.full((nq, 2), -1, dtype=np.int32)
    if nq == 0 or t.shape[0] == 0:
        return idx, dist
    bf = cv2.BFMatcher(cv2.NORM_HAMMING)
    res = bf.knnMatch(q, t, k=2)
    for i, ms in enumerate(res):
        for j, m in enumerate(ms[:2]):
            idx[i, j] = m.trainIdx
            dist[i, j] = int(m.distance)
    return idx, dist


def knn2_numpy(desc_to, desc_from):
    """Independent restatement of the same 2-NN (popcount table; lowest index on ties)."""
    q = np.ascontiguousarray(desc_to, dtype=np.uint8)
    t = np.ascontiguousarray(desc_from, dtype=np.uint8)
    nq, nt = q.shape[0], t.shape[0]
    idx = np.full((nq, 2), -1, dtype=np.int32)
    dist = np.full((nq, 2), -1, dtype=np.int32)
    if nq == 0 or nt == 0:
        return idx, dist
    pop = np.array([bin(i).count("1") for i in range(256)], dtype=np.int32)
    for s in range(0, nq, 256):
        x = np.bitwise_xor(q[s:s + 256, None, :], t[None, :, :])
        d = pop[x].sum(axis=2)
        o = np.argsort(d, axis=1, kind="stable")[:, :2]
        dv = np.take_along_axis(d, o, axis=1)
        idx[s:s + 256, :o.shape[1]] = o
        dist[s:s + 256, :o.shape[1]] = dv
    return idx, dist


def nndr_accept(d1, d2, nndr, strict=False):
    """f32 compare as in RTAB-Map: global path rejects iff d1 > nndr*d2 (accept d1 <= nndr*d2);
    guess-window path (myRegistrationVis.cpp:584,744) is strict d1 < nndr*d2."""
    a = np.float32(d1)
    b = np.float32(nndr) * np.float32(d2)
    return bool(a < b) if strict else bool(a <= b)


def global_match(desc_to, desc_from, nndr=0.8, strict=False, knn=knn2):
    """myRegistrationVis.cpp:839-894 -> list of (from_idx, to_idx) sorted by from_idx.

    To row j is assigned From id i1 iff it has two neighbours and passes NNDR; To ids that
    occur more than once are dropped on the To side (:882-894) -- both rows go."""
    idx, dist = knn(desc_to, desc_from)
    nq = idx.shape[0]
    assigned = np.full(nq, -1, dtype=np.int64)
    for j in range(nq):
        if idx[j, 1] < 0:
            continue
        if nndr_accept(dist[j, 0], dist[j, 1], nndr, strict):
            assigned[j] = idx[j, 0]
    counts = {}
    for j in range(nq):
        if assigned[j] >= 0:
            counts[int(assigned[j])] = counts.get(int(assigned[j]), 0) + 1
    pairs = [(int(assigned[j]), j) for j in range(nq) if assigned[j] >= 0 and counts[int(assigned[j])] == 1]
    pairs.sort()
    return pairs


def pnp_correspondences(pairs, kp3d_from, kp2d_to):
    """ascending From id; drop correspondences whose From 3D point is not finite.
    Returns (obj[M,3] f32, img[M,2] f32, from_ids[M], to_ids[M])."""
    obj, img, fi, ti = [], [], [], []
    for (f, t) in pairs:
        p = kp3d_from[f]
        if not np.all(np.isfinite(p)):
            continue
        obj.append(p)
        img.append(kp2d_to[t])
        fi.append(f)
        ti.append(t)
    return (np.asarray(obj, dtype=np.float32).reshape(-1, 3), np.asarray(img, dtype=np.float32).reshape(-1, 2),
            np.asarray(fi, dtype=np.int32), np.asarray(ti, dtype=np.int32))


# ---- pass 2: matching inside a window around the guess (R/src/myRegistrationVis.cpp:477-825) -------------
def _octave(o):
    """:707-708 / :725-726: SIFT-packed octave, low byte sign-extended."""
    o = int(o) & 255
    return o if o < 128 else (-128 | o)


def project_with_guess(kp3d_from, rvec, tvec, K, width, height):
    """:486-512: cv::projectPoints of the From 3-D points with guessCameraRef = (guess * localTransform)^-1, kept iff
    uIsInBounds(x, 0, width-1) and uIsInBounds(y, 0, height-1) (UtiLite: low <= v < high, NaN -> false) and z > 0.
    Returns (projected [N,2] f32, valid [N] bool)."""
    p3 = np.ascontiguousarray(kp3d_from, dtype=np.float32).reshape(-1, 1, 3)
    n = p3.shape[0]
    if n == 0:
        return np.zeros((0, 2), np.float32), np.zeros(0, bool)
    finite = np.isfinite(p3.reshape(-1, 3)).all(axis=1)
    safe = np.where(finite[:, None, None], p3, np.float32(1.0))
    proj, _ = cv2.projectPoints(safe, np.asarray(rvec, np.float64).reshape(3, 1), np.asarray(tvec, np.float64).reshape(3, 1),
                                np.asarray(K, np.float64), None)
    proj = proj.reshape(-1, 2).astype(np.float32)
    R, _ = cv2.Rodrigues(np.asarray(rvec, np.float64).reshape(3, 1))
    z = p3.reshape(-1, 3).astype(np.float64) @ R[2] + float(np.asarray(tvec).ravel()[2])
    x, y = proj[:, 0], proj[:, 1]
    valid = finite & (x >= np.float32(0)) & (x < np.float32(width - 1)) & (y >= np.float32(0)) & (y < np.float32(height - 1)) & (z > 0)
    return proj, valid


def guess_match(desc_from, desc_to, kp3d_from, kp2d_to, oct_from, oct_to, rvec, tvec, K, width, height,
                win=20.0, nndr=0.8):
    """Default branch (:667-818, Vis/CorGuessMatchToProjection=false) with an EXACT radius search standing in for
    FLANN's approximate randomised kd-tree (:672-680; checks=32, not reproducible): candidates of a projected From
    feature = To keypoints with squared pixel distance < win^2 (L2_Simple: dx*dx + dy*dy in f32), ordered by
    (distance, index) like a sorted result set.  >= 2 candidates: keep those with the From feature's octave; >= 2 left ->
    BFMatcher(HAMMING).knnMatch(k=2) among them (ties -> first in candidate order), accept iff d1 < nndr*d2 (strict,
    :744); exactly 1 left -> accept it (:751-754).  1 candidate: accept iff octaves agree (:756-764).  A To feature
    belongs to the first From feature (ascending index) that claims it (:776-787).
    Returns [(from_idx, to_idx)] ascending in from_idx."""
    desc_from = np.ascontiguousarray(desc_from, np.uint8)
    desc_to = np.ascontiguousarray(desc_to, np.uint8)
    kp2d_to = np.ascontiguousarray(kp2d_to, np.float32).reshape(-1, 2)
    proj, valid = project_with_guess(kp3d_from, rvec, tvec, K, width, height)
    r2 = np.float32(win) * np.float32(win)
    pop = np.array([bin(i).count("1") for i in range(256)], dtype=np.int32)
    taken, pairs = set(), []
    for i in np.nonzero(valid)[0]:
        dx = proj[i, 0] - kp2d_to[:, 0]
        dy = proj[i, 1] - kp2d_to[:, 1]
        d2 = (dx * dx).astype(np.float32) + (dy * dy).astype(np.float32)
        cand = np.nonzero(d2 < r2)[0]
        cand = cand[np.lexsort((cand, d2[cand]))]
        of = _octave(oct_from[i])
        matched = -1
        if len(cand) >= 2:
            filt = [int(j) for j in cand if _octave(oct_to[j]) == of]
            if len(filt) >= 2:
                ham = pop[np.bitwise_xor(desc_to[filt], desc_from[i][None, :])].sum(axis=1)
                o = np.argsort(ham, kind="stable")
                if nndr_accept(ham[o[0]], ham[o[1]], nndr, strict=True):
                    matched = filt[int(o[0])]
            elif len(filt) == 1:
                matched = filt[0]
        elif len(cand) == 1:
            if _octave(oct_to[cand[0]]) == of:
                matched = int(cand[0])
        if matched >= 0 and matched not in taken:
            taken.add(matched)
            pairs.append((int(i), matched))
    return pairs


def guess_from_transform(T3x4, local_R, local_t):
    """guessCameraRef = (guess * localTransform)^-1 (:486-488) from the f32 3x4 transform `guess` (From-base <- To-base)
    and the camera's localTransform.  rtabmap::Transform holds f32: the product and the inverse's translation are
    rounded to f32 (the convention of sepb_pnp_set_guess); Rodrigues as cv::Rodrigues -> (rvec [3], tvec [3])."""
    T = np.asarray(T3x4, np.float32).reshape(3, 4).astype(np.float64)
    Rl, tl = np.asarray(local_R, np.float64).reshape(3, 3), np.asarray(local_t, np.float64).reshape(3)
    RP = np.empty((3, 3))
    tP = np.empty(3)
    for i in range(3):
        for j in range(3):
            RP[i, j] = np.float32(T[i, 0] * Rl[0, j] + T[i, 1] * Rl[1, j] + T[i, 2] * Rl[2, j])
        tP[i] = np.float32(T[i, 0] * tl[0] + T[i, 1] * tl[1] + T[i, 2] * tl[2] + T[i, 3])
    Rg = RP.T.copy()
    tg = np.array([np.float32(-(Rg[i, 0] * tP[0] + Rg[i, 1] * tP[1] + Rg[i, 2] * tP[2])) for i in range(3)], np.float64)
    rvec, _ = cv2.Rodrigues(Rg)
    return rvec.ravel(), tg


def transform_from_pnp(rvec, tvec, local_R, local_t):
    """util3d::estimateMotion3DTo2D: T = (localTransform * [R|t])^-1 as the f32 3x4 rtabmap::Transform (row-major 12)."""
    R, _ = cv2.Rodrigues(np.asarray(rvec, np.float64).reshape(3, 1))
    Rl, tl = np.asarray(local_R, np.float64).reshape(3, 3), np.asarray(local_t, np.float64).reshape(3)
    RP = Rl @ R
    tP = Rl @ np.asarray(tvec, np.float64).reshape(3) + tl
    T = np.empty((3, 4), np.float32)
    T[:, :3] = RP.T
    T[:, 3] = -(RP.T @ tP)
    return T.reshape(12)
