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
