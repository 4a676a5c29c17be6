"""Batch driver of the separator-detection hot path over the C ABI.

Reproduces the reference's call order for one query round:
`find_matches` (NetVLAD NN, R/scripts/data_handler.py:166-209) on the computing robot, then for
every candidate pair `estimate_transformation` pass 1 (R/src/stereoCamGeometricTools.cpp:122-152)
with descriptorsFrom = querying robot's frame, descriptorsTo = computing robot's frame
(R/scripts/find_separators.py:85-91) -- batched: all candidate pairs of a round go through
K1 -> K2 -> K3 on the device without host round trips.

torch is used only for device buffers / streams (plumbing); all compute is libsepb200.so.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np
import torch

from . import capi


@dataclass
class Camera:
    fx: float
    fy: float
    cx: float
    cy: float


class SeparatorEngine:
    """One computing robot (local frames = To side, distance-matrix rows) against one querying
    peer (received NetVLAD descriptors = database columns, its frames = From side)."""

    def __init__(self, cam: Camera, max_feat: int, dim: int, local_capacity: int, peer_capacity: int, device: int = 0,
                 iterations: int = 300, min_inliers: int = 5, reproj_error: float = 2.0, nndr: float = 0.8,
                 db_base: int = 0, ctx: capi.Context | None = None, image_size=(1241, 376), guess_win_size: float = 20.0,
                 pnp_flags: int = 0, refine_iterations: int = 0, minimal_model: bool = False,
                 search_ctx: capi.Context | None = None):
        """ctx: share another engine's context (one stream, one copy stream): two engines then act as the two
        buffers of a pipeline -- the next round's keyframes upload while the current round is verified."""
        self.device = device
        torch.cuda.set_device(device)
        self._owns_ctx = ctx is None
        self.ctx = capi.Context(device) if ctx is None else ctx
        self.stream = torch.cuda.ExternalStream(self.ctx.stream, device=device)
        self.max_feat, self.dim = max_feat, dim
        self.local = capi.Frames(self.ctx, local_capacity, max_feat)
        self.peer = capi.Frames(self.ctx, peer_capacity, max_feat)
        # search_ctx: the NetVLAD database lives on another context of the same device (its own, typically high-priority,
        # stream), so the candidate search of the next round runs while this context verifies the current one
        self.search_ctx = search_ctx if search_ctx is not None else self.ctx
        self.db = capi.NetvladDB(self.search_ctx, dim, peer_capacity, base=db_base)
        self.db_base = db_base
        self.pnp = capi.pnp_params(cam.fx, cam.fy, cam.cx, cam.cy, iterations=iterations, min_inliers=min_inliers,
                                   reproj_error=reproj_error, flags=pnp_flags, refine_iterations=refine_iterations,
                                   minimal_model=minimal_model)
        self.mp = capi.MatchParams(nndr, 0)
        self.guess = capi.guess_params(image_size[0], image_size[1], guess_win_size)
        self.n_local = 0
        self.n_peer = 0
        self._buf = {}

    # ---- data in (host buffers; H2D inside the C ABI) -------------------------------------
    def add_local_keyframes(self, desc, kp2d, kp3d, count, octave=None, async_=False, lazy=False):
        """lazy=True: host-resident keyframes (pinned arrays): descriptors travel now, key points are pulled by the device
        for the frames whose pairs reach PnP (capi.Frames.upload)."""
        self.local.upload(self.n_local, desc, kp2d, kp3d, count, octave, async_=async_, lazy=lazy)
        self.n_local += len(count)

    def add_peer_keyframes(self, desc, kp2d, kp3d, count, octave=None, async_=False, lazy=False):
        self.peer.upload(self.n_peer, desc, kp2d, kp3d, count, octave, async_=async_, lazy=lazy)
        self.n_peer += len(count)

    def receive_descriptors(self, netvlad_rows):
        """find_matches_service request payload (float64[] reshaped (-1, dim), data_handler.py:300-301)."""
        self.db.append(netvlad_rows)

    def reset(self):
        """Forget all keyframes / descriptors (device buffers are kept and overwritten)."""
        self.n_local = self.n_peer = 0
        self.db.clear()

    # ---- device-resident step ----------------------------------------------------------------
    def _tensor(self, name, shape, dtype):
        t = self._buf.get(name)
        if t is None or tuple(t.shape) != tuple(shape) or t.dtype != dtype:
            t = torch.empty(shape, dtype=dtype, device=f"cuda:{self.device}")
            self._buf[name] = t
        return t

    def candidates_dev(self, d_queries: torch.Tensor, k: int, mode=capi.NN_EXACT, row_mask=None):
        """K1 on resident queries [Q, dim] f64 -> (idx [Q,k] i32, dist [Q,k] f64) device tensors."""
        Q = d_queries.shape[0]
        idx = self._tensor("nn_idx", (Q, k), torch.int32)
        dist = self._tensor("nn_dist", (Q, k), torch.float64)
        rm = C.c_void_p(row_mask.data_ptr()) if row_mask is not None else None
        self.search_ctx.check(capi.lib().sepb_netvlad_topk_dev(self.search_ctx._h, self.db._h, C.c_void_p(d_queries.data_ptr()), Q, k,
                                                        mode, rm, None, None, C.c_void_p(idx.data_ptr()),
                                                        C.c_void_p(dist.data_ptr())))
        if self.search_ctx is not self.ctx:
            self.search_ctx.sync()      # the verification stream reads idx next
        return idx, dist

    def pairs_from_candidates(self, idx: torch.Tensor, base: int | None = None):
        """(from = peer frame slot = database row - base, to = local frame = query row); -1 = no candidate.
        Database rows are GLOBAL ids (`sepb_db_set_base`): the frame store of this engine holds the rows of its own
        shard, so the shard's base is subtracted; rows of other shards (after a merge) map to -1 here."""
        Q, k = idx.shape
        base = self.db_base if base is None else base
        with torch.cuda.stream(self.stream):
            rows = torch.arange(Q, dtype=torch.int32, device=idx.device).repeat_interleave(k)
            col = idx.reshape(-1)
            if base:
                local = col - base
                col = torch.where((col >= 0) & (local >= 0) & (local < self.peer.capacity), local, torch.full_like(col, -1))
            pairs = torch.stack([col, rows], dim=1).contiguous()
        return pairs

    def verify_dev(self, d_pairs: torch.Tensor, two_pass: bool = False):
        """K2 -> K3 on resident pairs [P,2] i32 -> dict of device tensors (asynchronous).
        two_pass: the whole estimate_transformation call (second pass with guess = first result, K2G -> K3)."""
        P, F = d_pairs.shape[0], self.max_feat
        o = dict(status=self._tensor("status", (P,), torch.int32), n_matches=self._tensor("n_matches", (P,), torch.int32),
                 n_inliers=self._tensor("n_inliers", (P,), torch.int32), pose=self._tensor("pose", (P, 6), torch.float64),
                 transform=self._tensor("transform", (P, 12), torch.float32),
                 match_from=self._tensor("match_from", (P, F), torch.int16),
                 match_to=self._tensor("match_to", (P, F), torch.int16),
                 inlier_mask=self._tensor("inlier_mask", (P, (F + 31) // 32), torch.int32),
                 best_hyp=self._tensor("best_hyp", (P,), torch.int32), iters_run=self._tensor("iters_run", (P,), torch.int32),
                 covariance=self._tensor("covariance", (P, 36), torch.float64))
        out = capi.VerifyOut(**{k: v.data_ptr() for k, v in o.items()})
        if two_pass:
            st1 = self._tensor("status_pass1", (P,), torch.int32)
            self.ctx.check(capi.lib().sepb_estimate_transformations_dev(
                self.ctx._h, self.peer._h, self.local._h, C.c_void_p(d_pairs.data_ptr()), P, C.byref(self.mp),
                C.byref(self.pnp), C.byref(self.guess), C.byref(out), C.c_void_p(st1.data_ptr())))
            o = dict(o, status_pass1=st1)
        else:
            self.ctx.check(capi.lib().sepb_verify_pairs_dev(self.ctx._h, self.peer._h, self.local._h,
                                                            C.c_void_p(d_pairs.data_ptr()), P, C.byref(self.mp),
                                                            C.byref(self.pnp), None, 0, C.byref(out)))
        return o

    def step_dev(self, d_queries: torch.Tensor, k: int, mode=capi.NN_EXACT):
        """One pass of the hot path with everything resident: K1 top-k, pair list, K2, K3."""
        idx, dist = self.candidates_dev(d_queries, k, mode)
        pairs = self.pairs_from_candidates(idx)
        res = self.verify_dev(pairs)
        res["pairs"], res["nn_dist"] = pairs, dist
        return res

    # ---- host-buffer round (what a ROS-side caller does) -------------------------------------
    def find_matches(self, local_descriptors, netvlad_distance=0.13, max_matches_nb=20, local_used=(), other_used=(),
                     pairs_ignored=(), mode=capi.NN_EXACT):
        """DataHandler.find_matches (data_handler.py:166-209): list of (idx_local, idx_other)."""
        self.db.set_col_mask(other_used)
        i1, d1 = self.db.topk(local_descriptors, 1, mode, local_used, pairs_ignored)
        return capi.select_matches(i1, d1, netvlad_distance, max_matches_nb)

    def estimate_transformations(self, matches, two_pass: bool = False):
        """estimate_transformation for [(idx_local, idx_other)]: From = peer frame, To = local frame.  two_pass=False:
        the first call only (global matching, stereoCamGeometricTools.cpp:141-152); True: both calls (:141-164)."""
        pairs = np.array([[o, l] for (l, o) in matches], dtype=np.int32).reshape(-1, 2)
        if two_pass:
            return capi.estimate_transformations(self.ctx, self.peer, self.local, pairs, self.pnp, self.guess, self.mp.nndr)
        return capi.verify_pairs(self.ctx, self.peer, self.local, pairs, self.pnp, self.mp.nndr, bool(self.mp.nndr_strict))

    def sync(self):
        self.ctx.sync()

    def close(self):
        for obj in (self.local, self.peer, self.db) + ((self.ctx,) if self._owns_ctx else ()):
            obj.close()
