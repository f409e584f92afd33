"""Data-parallel sharding of a pair batch over the GPUs of one box.

Image pairs are independent along the whole path (the reference never mixes the batch dimension: encoder per image,
attention per pair, convolutions per image, refinement per pixel), so N GPUs run N engine replicas on contiguous
batch shards with NO collective on the data path.  The only communication is the optional gather of the finished
outputs (2.8 MB / pair) after the forward: ``torch.distributed.all_gather`` (NCCL over NVLink on the GPU box, gloo in
the CPU tests).  One process per GPU (torchrun); each rank constructs the model on its own ``cuda:LOCAL_RANK``.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Tuple

import torch
import torch.distributed as dist


def shard_range(n_pairs: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous shard [lo, hi) of rank ``rank``: ceil(B / G) pairs per rank, trailing ranks may be empty (B < G leaves
    GPUs idle -- reported, never padded)."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"bad rank {rank} / world {world}")
    per = (n_pairs + world - 1) // world
    lo = min(rank * per, n_pairs)
    return lo, min(lo + per, n_pairs)


def shard_sizes(n_pairs: int, world: int) -> List[int]:
    return [hi - lo for lo, hi in (shard_range(n_pairs, world, r) for r in range(world))]


def _gather_ragged(local: torch.Tensor, sizes: List[int], group) -> torch.Tensor:
    """All-gather tensors whose leading dimension differs per rank into ONE [sum(sizes), ...] tensor.

    Equal shards (B divisible by the world size: the normal case) are gathered by ``all_gather_into_tensor`` straight
    into the final buffer -- no padding, no per-rank staging buffers, no concatenation.  Ragged shards pad the local
    shard to the largest one for the collective and compact the result once."""
    per = max(sizes)
    world = len(sizes)
    tail = tuple(local.shape[1:])
    local = local.contiguous()
    if local.shape[0] != per:
        pad = torch.zeros((per,) + tail, dtype=local.dtype, device=local.device)
        pad[: local.shape[0]] = local
        local = pad
    out = torch.empty((world * per,) + tail, dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, local, group=group)
    if all(s == per for s in sizes):
        return out
    return torch.cat([out[r * per: r * per + s] for r, s in enumerate(sizes)], dim=0)


def predict_sharded(predict: Callable[[torch.Tensor, torch.Tensor], Tuple[torch.Tensor, Optional[torch.Tensor]]],
                    source: torch.Tensor, target: torch.Tensor, gather: bool = True, group=None):
    """Run ``predict(src_shard, tgt_shard) -> (flow [b,2,H,W], covisibility [b,H,W] | None)`` on this rank's contiguous
    shard of the batch and, if ``gather``, return the full-batch outputs on every rank (else the local shard).

    ``predict`` is normally ``lambda s, t: (o := model.predict_correspondences_batched(s, t)).flow.flow_output, ...``;
    every rank must be called with the same full batch (or at least the same batch size)."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    B = source.shape[0]
    lo, hi = shard_range(B, world, rank)
    flow = cov = None
    if hi > lo:
        flow, cov = predict(source[lo:hi], target[lo:hi])
    if not gather or world == 1:
        return flow, cov
    sizes = shard_sizes(B, world)
    # ranks with an empty shard still take part in the collective; shapes come from rank 0 (always non-empty)
    meta = [None]
    if rank == 0:
        meta = [(tuple(flow.shape[1:]), flow.dtype, None if cov is None else tuple(cov.shape[1:]))]
    dist.broadcast_object_list(meta, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    fshape, dtype, cshape = meta[0]
    if flow is not None:
        dev = flow.device
    elif dist.get_backend(group) == "nccl":
        # an empty shard (B < world size) still joins the collective: NCCL needs its buffers on this rank's GPU even when
        # the inputs are host tensors
        dev = torch.device("cuda", torch.cuda.current_device())
    else:
        dev = source.device
    if flow is None:
        flow = torch.zeros((0,) + fshape, dtype=dtype, device=dev)
        cov = None if cshape is None else torch.zeros((0,) + cshape, dtype=dtype, device=dev)
    out_flow = _gather_ragged(flow, sizes, group)
    out_cov = None if cshape is None else _gather_ragged(cov, sizes, group)
    return out_flow, out_cov
