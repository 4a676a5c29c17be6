"""Multi-GPU host logic of the hot path, one process per GPU over torch.distributed.

Two independent ways the path shards:

* verification pairs are independent units: `pair_block` gives every rank a contiguous block of
  the candidate list, no collective on the data path (results are concatenated by the caller);
* the peer NetVLAD database shards by rows: `shard_rows` gives rank r the rows
  [lo, hi); every rank searches its shard with GLOBAL row ids (`sepb_db_set_base`), the per-rank
  `[Q, k]` (idx, exact f64 dist) lists are exchanged with ONE all-gather (the only collective of
  the path, 12 bytes per candidate) and merged with the same (distance, index) order the
  single-GPU search uses (`sepb_topk_merge_dev`), so the result is identical to one big database.

torch.distributed is plumbing here (NCCL on GPUs, gloo in the CPU tests); the merge itself is
passed in, so the CPU tests can exercise the exchange layout without a CUDA device.
"""
from __future__ import annotations

import ctypes as C
from typing import Callable, Tuple

import torch
import torch.distributed as dist


def shard_rows(n_rows: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous row range [lo, hi) of a database of n_rows owned by `rank` (sizes differ by <= 1)."""
    if world <= 0 or not (0 <= rank < world) or n_rows < 0:
        raise ValueError("bad shard request")
    q, r = divmod(n_rows, world)
    lo = rank * q + min(rank, r)
    return lo, lo + q + (1 if rank < r else 0)


def pair_block(n_pairs: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block of the candidate pair list verified by `rank`."""
    return shard_rows(n_pairs, rank, world)


def gather_topk(idx: torch.Tensor, dist_: torch.Tensor, group=None) -> Tuple[torch.Tensor, torch.Tensor]:
    """All-gather the per-shard top-k lists: [Q, k] -> [world, Q, k] (layout sepb_topk_merge_dev takes)."""
    world = dist.get_world_size(group)
    Q, k = idx.shape
    all_idx = torch.empty((world, Q, k), dtype=idx.dtype, device=idx.device)
    all_dist = torch.empty((world, Q, k), dtype=dist_.dtype, device=dist_.device)
    dist.all_gather_into_tensor(all_idx.view(-1), idx.contiguous().view(-1), group=group)
    dist.all_gather_into_tensor(all_dist.view(-1), dist_.contiguous().view(-1), group=group)
    return all_idx, all_dist


def _order(after: "torch.cuda.Stream", before: "torch.cuda.Stream"):
    """Work enqueued on `after` from now on waits for everything enqueued on `before` so far (device-side)."""
    if after.cuda_stream != before.cuda_stream:
        ev = torch.cuda.Event()
        ev.record(before)
        after.wait_event(ev)


def merge_on_device(ctx) -> Callable[[torch.Tensor, torch.Tensor], Tuple[torch.Tensor, torch.Tensor]]:
    """The product merge: sepb_topk_merge_dev on the context's stream (fails loudly without the library).

    The library's stream is non-blocking and the collective runs on torch's current stream, so the two are ordered
    explicitly, both ways: the merge waits for the gather, and torch's current stream then waits for the merge --
    the returned tensors are safe to use on the current stream like any torch result."""
    from . import capi

    def merge(all_idx: torch.Tensor, all_dist: torch.Tensor):
        lists, Q, k = all_idx.shape
        out_i = torch.empty((Q, k), dtype=torch.int32, device=all_idx.device)
        out_d = torch.empty((Q, k), dtype=torch.float64, device=all_idx.device)
        lib_stream = torch.cuda.ExternalStream(ctx.stream, device=all_idx.device)
        cur = torch.cuda.current_stream(all_idx.device)
        _order(lib_stream, cur)
        ctx.check(capi.lib().sepb_topk_merge_dev(ctx._h, C.c_void_p(all_idx.data_ptr()), C.c_void_p(all_dist.data_ptr()),
                                                 lists, Q, k, C.c_void_p(out_i.data_ptr()), C.c_void_p(out_d.data_ptr())))
        # torch's stream now waits for the merge: the caching allocator may reuse these buffers on it safely (no
        # record_stream on the library's stream -- it can be destroyed before the tensors are)
        _order(cur, lib_stream)
        return out_i, out_d

    return merge


def sharded_topk(local_topk: Callable[[], Tuple[torch.Tensor, torch.Tensor]], merge, group=None, sync=None, ctx=None):
    """One sharded NetVLAD search: local exact top-k with global ids -> all-gather -> merge.

    local_topk() returns this rank's (idx [Q,k] i32, dist [Q,k] f64) with GLOBAL row ids.  ctx: the library context
    local_topk enqueued on -- the collective (torch's current stream) is then ordered behind it on the device;
    `sync` (optional, legacy) is a host-side alternative."""
    idx, d = local_topk()
    if ctx is not None and idx.is_cuda:
        _order(torch.cuda.current_stream(idx.device), torch.cuda.ExternalStream(ctx.stream, device=idx.device))
    if sync is not None:
        sync()
    all_idx, all_d = gather_topk(idx, d, group)
    return merge(all_idx, all_d)
