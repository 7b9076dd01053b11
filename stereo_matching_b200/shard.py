"""Batch partition of independent stereo pairs over ranks.

The reference processes pairs one after another in a Python loop
(/root/reference/inference.py:99) with no state carried between them, so a
multi-GPU run shards the pair list and needs no data-path collective; only the
per-pair results (uint8 maps, 3-px errors) are gathered on rank 0.
"""
from __future__ import annotations

from typing import List, Sequence


def shard_indices(n_items: int, rank: int, world_size: int) -> List[int]:
    """Round-robin shard: item i goes to rank i % world_size (keeps ranks within
    one item of each other for any n_items)."""
    if world_size < 1 or not 0 <= rank < world_size:
        raise ValueError(f"bad rank {rank} / world size {world_size}")
    return list(range(rank, n_items, world_size))


def gather_results(local: Sequence, n_items: int, rank: int, world_size: int, group=None):
    """Collect per-item results of `shard_indices` order back into item order on
    every rank (torch.distributed all_gather_object; gloo on CPU, NCCL needs none
    of this on the data path)."""
    if world_size == 1:
        return list(local)
    import torch.distributed as dist
    gathered = [None] * world_size
    dist.all_gather_object(gathered, list(local), group=group)
    out = [None] * n_items
    for r, items in enumerate(gathered):
        for k, i in enumerate(shard_indices(n_items, r, world_size)):
            out[i] = items[k]
    return out


def mean_error(errors_local: Sequence[float], n_items: int, rank: int, world_size: int, group=None) -> float:
    """`error / FLAGS.num_val_img` of inference.py:236 over all ranks' pairs."""
    all_err = gather_results(errors_local, n_items, rank, world_size, group)
    return float(sum(all_err) / max(n_items, 1))
