"""Row-sharded database over torch.distributed (one process per GPU).

The reference has no multi-device code; its per-row
independent encode (product_quantization.py:76-80) and its flat search
(flat_quantized_index.py:45-76) shard by database rows:

  * rank g owns rows [offset_g, offset_g + n_g); encode needs no collective;
  * search: every rank scans its shard for a local top-k, the (nq, k)
    (value, global id) pairs are exchanged with ONE all-gather (NCCL over
    NVLink on the GPU box, gloo in the CPU tests) and every rank runs the same
    deterministic k-way merge ordered by (value, id);
  * training statistics (Lloyd sums / counts, min / max) are all-reduced.

`merge_fn` is the device merge kernel (ops.topk_merge) on the GPU; the CPU
tests pass a numpy merge so the exchange logic is covered with gloo.
"""
from __future__ import annotations

from typing import Callable, Optional, Tuple

import torch
import torch.distributed as dist


def shard_bounds(n_total: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous row range of `rank`: the first n_total % world ranks get one extra row."""
    base, rem = divmod(int(n_total), int(world))
    start = rank * base + min(rank, rem)
    return start, start + base + (1 if rank < rem else 0)


def _world(group=None) -> Tuple[int, int]:
    if dist.is_available() and dist.is_initialized():
        return dist.get_world_size(group), dist.get_rank(group)
    return 1, 0


def gather_topk(vals: torch.Tensor, ids_local: torch.Tensor, row_offset: int, group=None
                ) -> Tuple[torch.Tensor, torch.Tensor]:
    """All-gather per-shard top-k lists.

    vals (nq, k) fp32, ids_local (nq, k) int32 holding uint32 local row ids
    (0xffffffff = empty slot). Returns (nq, P, k) fp32 and (nq, P, k) int64
    GLOBAL ids (-1 = empty) with P = world size, identical on every rank.
    """
    world, _ = _world(group)
    ids64 = ids_local.to(torch.int64) & 0xFFFFFFFF
    empty = ids64 == 0xFFFFFFFF
    ids64 = torch.where(empty, torch.full_like(ids64, -1), ids64 + int(row_offset))
    if world == 1:
        return vals.unsqueeze(1).contiguous(), ids64.unsqueeze(1).contiguous()
    nq, k = vals.shape
    # (world * nq, k) concatenation along dim 0: the form both NCCL and gloo accept
    gv = torch.empty((world * nq, k), dtype=vals.dtype, device=vals.device)
    gi = torch.empty((world * nq, k), dtype=torch.int64, device=vals.device)
    dist.all_gather_into_tensor(gv, vals.contiguous(), group=group)
    dist.all_gather_into_tensor(gi, ids64.contiguous(), group=group)
    return (gv.view(world, nq, k).permute(1, 0, 2).contiguous(),
            gi.view(world, nq, k).permute(1, 0, 2).contiguous())


def sharded_topk(vals: torch.Tensor, ids_local: torch.Tensor, row_offset: int, k: int, metric,
                 merge_fn: Optional[Callable] = None, group=None):
    """Local (nq, k) lists -> global (nq, k) (vals fp32, ids int64), same on every rank."""
    gv, gi = gather_topk(vals, ids_local, row_offset, group)
    if merge_fn is None:
        from . import ops
        merge_fn = ops.topk_merge
    return merge_fn(gv, gi, k, metric)


def allreduce_kmeans_stats(sums: torch.Tensor, counts: torch.Tensor, group=None) -> None:
    """Sum Lloyd accumulators over the shards (in place)."""
    world, _ = _world(group)
    if world > 1:
        dist.all_reduce(sums, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(counts, op=dist.ReduceOp.SUM, group=group)


def allreduce_minmax(vmin: torch.Tensor, vmax: torch.Tensor, group=None) -> None:
    world, _ = _world(group)
    if world > 1:
        dist.all_reduce(vmin, op=dist.ReduceOp.MIN, group=group)
        dist.all_reduce(vmax, op=dist.ReduceOp.MAX, group=group)


def numpy_merge(gv: torch.Tensor, gi: torch.Tensor, k: int, metric):
    """Host merge with the kernel's ordering ((value, id) ascending; IP by descending value).
    Only used by the CPU (gloo) tests of the exchange logic."""
    import numpy as np
    v = gv.cpu().numpy().reshape(gv.shape[0], -1)
    i = gi.cpu().numpy().reshape(gi.shape[0], -1)
    ip = metric in ("ip", "IP", 1)
    key = np.where(i < 0, np.inf, -v if ip else v)
    ids_key = np.where(i < 0, np.iinfo(np.int64).max, i)
    order = np.lexsort((ids_key, key), axis=1)[:, :k]
    ov = np.take_along_axis(v, order, 1)
    oi = np.take_along_axis(i, order, 1)
    bad = np.take_along_axis(i, order, 1) < 0
    ov = np.where(bad, -np.inf if ip else np.inf, ov).astype(np.float32)
    return torch.from_numpy(ov), torch.from_numpy(oi)
