"""Batch sharding of the hot path over the GPUs of one box.

Sequences are independent and every neuron state is private to its sample, so inference shards by
sample with NO data-path collective: rank r bins and forwards its own contiguous slice of the batch
(the reference trains with DDP, config/config.yaml:35, i.e. the same per-sample split).  The only
collective on this path is the optional gather of the per-sample predictions for reporting, and - in
training - the flat gradient all-reduce.  Both go through ``torch.distributed`` (NCCL on GPUs; the CPU
tests run the same code over ``gloo``).
"""

from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

__all__ = ["shard_range", "shard_samples", "gather_predictions", "allreduce_flat_gradients", "convert_sync_batchnorm"]


def shard_range(n: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous ``[lo, hi)`` slice of ``n`` samples owned by ``rank``; sizes differ by at most one."""
    if not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside world of {world}")
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_samples(samples: Sequence, rank: Optional[int] = None, world: Optional[int] = None) -> List:
    """This rank's event streams (list of per-sample record arrays)."""
    if rank is None:
        rank = dist.get_rank() if dist.is_initialized() else 0
    if world is None:
        world = dist.get_world_size() if dist.is_initialized() else 1
    lo, hi = shard_range(len(samples), rank, world)
    return list(samples[lo:hi])


def gather_predictions(cls: torch.Tensor, bbox: torch.Tensor, n_total: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """All-gather of the ragged per-rank ``cls [b_r, A, C]`` / ``bbox [b_r, A, 4]`` into batch order."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return cls, bbox
    world = dist.get_world_size()
    sizes = [shard_range(n_total, r, world) for r in range(world)]
    bmax = max(hi - lo for lo, hi in sizes)
    outs = []
    for t in (cls, bbox):
        pad = torch.zeros((bmax,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        pad[: t.shape[0]] = t
        buf = [torch.empty_like(pad) for _ in range(world)]
        dist.all_gather(buf, pad)
        outs.append(torch.cat([b[: hi - lo] for b, (lo, hi) in zip(buf, sizes)]))
    return outs[0], outs[1]


def allreduce_flat_gradients(params: Sequence[torch.nn.Parameter]) -> None:
    """One all-reduce of the flattened gradient (mean over ranks), written back in place - the only
    training collective (4,228,544 fp32 = 16.9 MB for TinyYolo)."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return
    grads = [p.grad for p in params if p.grad is not None]
    if not grads:
        return
    flat = torch.cat([g.reshape(-1) for g in grads])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM)
    flat /= dist.get_world_size()
    off = 0
    for g in grads:
        n = g.numel()
        g.copy_(flat[off : off + n].view_as(g))
        off += n


def convert_sync_batchnorm(model: torch.nn.Module) -> torch.nn.Module:
    """Role of ``nn.SyncBatchNorm.convert_sync_batchnorm`` / Lightning's ``sync_batchnorm: true`` (reference
    ``config.yaml:76``): every train-mode BatchNorm takes its per-time-step statistics over the GLOBAL batch - one
    all-reduce of the ``[T, 2, C]`` sums per layer and direction (``soda_bn_train_stats`` / ``*_from_sums``)."""
    from .modules.common import SodaBatchNorm2d

    for m in model.modules():
        if isinstance(m, SodaBatchNorm2d):
            m.sync = True
    return model
