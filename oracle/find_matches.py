"""TEST INFRASTRUCTURE ONLY (oracle) -- never imported by the product path.

CPU restatement of the reference's NetVLAD nearest-neighbour search and candidate
selection, `DataHandler.find_matches` (R/scripts/data_handler.py:166-209), using the same
library call the reference makes (`scipy.spatial.distance.cdist`, f64 Euclidean, :170).

Tie rule: `np.argsort` (default quicksort) is unstable on exact ties (:187,:191); the pinned
rule here and in the CUDA path is "lowest index wins" (== `np.argmin` / stable argsort).
"""
from __future__ import annotations

import numpy as np
from scipy.spatial.distance import cdist


def distance_matrix(local_descs, received_descs, local_used=(), other_used=(), pairs_ignored=()):
    """data_handler.py:168-184: cdist + row / column / cell masks set to +inf."""
    d = cdist(np.asarray(local_descs, dtype=np.float64), np.asarray(received_descs, dtype=np.float64))
    if len(local_used) > 0:
        d[np.asarray(local_used, dtype=np.int64)] = np.inf
    if len(other_used) > 0:
        d[:, np.asarray(other_used, dtype=np.int64)] = np.inf
    for (a, b) in pairs_ignored:
        d[a, b] = np.inf
    return d


def select_matches(row_min_idx, row_min_val, netvlad_distance, max_matches_nb):
    """data_handler.py:191-205: global order of row minima, greedy accept.

    A row whose column was already taken is skipped but still consumes one of the
    `max_matches_nb` slots (:194,:199-200); the loop breaks at the first d >= threshold (:202-205).
    """
    order = np.argsort(row_min_val, kind="stable")
    matches = []
    for i in range(min(len(order), int(max_matches_nb))):
        idx_local = int(order[i])
        idx_other = int(row_min_idx[idx_local])
        if idx_other in [m[1] for m in matches]:
            continue
        if row_min_val[idx_local] < netvlad_distance:
            matches.append((idx_local, idx_other))
        else:
            break
    return matches


def find_matches(local_descs, received_descs, netvlad_distance=0.13, max_matches_nb=20,
                 local_used=(), other_used=(), pairs_ignored=()):
    """Full restatement of data_handler.py:166-209 -> list of (idx_local, idx_other)."""
    d = distance_matrix(local_descs, received_descs, local_used, other_used, pairs_ignored)
    row_min_idx = np.argmin(d, axis=1)  # == argsort(axis=1)[:, 0] with lowest-index ties
    row_min_val = d[np.arange(d.shape[0]), row_min_idx]
    return select_matches(row_min_idx, row_min_val, netvlad_distance, max_matches_nb)


def topk(local_descs, received_descs, k, local_used=(), other_used=(), pairs_ignored=(), chunk=1024):
    """Batch extension (BASELINE.json configs[1]: top-5): per local row the k smallest cdist
    distances, ascending, ties -> lowest column.  Returns (idx[Q,k] int32, dist[Q,k] f64);
    masked / missing entries are (-1, +inf).  Rows are processed in chunks to bound memory."""
    a = np.asarray(local_descs, dtype=np.float64)
    b = np.asarray(received_descs, dtype=np.float64)
    Q, N = a.shape[0], b.shape[0]
    idx = np.full((Q, k), -1, dtype=np.int32)
    dist = np.full((Q, k), np.inf, dtype=np.float64)
    used_rows = np.zeros(Q, dtype=bool)
    if len(local_used):
        used_rows[np.asarray(local_used, dtype=np.int64)] = True
    cell = {}
    for (r, c) in pairs_ignored:
        cell.setdefault(int(r), []).append(int(c))
    for s in range(0, Q, chunk):
        e = min(Q, s + chunk)
        d = cdist(a[s:e], b)
        if len(other_used):
            d[:, np.asarray(other_used, dtype=np.int64)] = np.inf
        d[used_rows[s:e]] = np.inf
        for r in range(s, e):
            if r in cell:
                d[r - s, cell[r]] = np.inf
        kk = min(k, N)
        o = np.argsort(d, axis=1, kind="stable")[:, :kk]
        dv = np.take_along_axis(d, o, axis=1)
        o = np.where(np.isinf(dv), -1, o)
        idx[s:e, :kk] = o
        dist[s:e, :kk] = dv
    return idx, dist
