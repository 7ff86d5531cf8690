"""Batch sharding across the GPUs of one box.

Configurations are independent units: rank r checks the contiguous shard
``[r * ceil(N / R), ...)`` on its own GPU with its own replica of the (tiny) robot / scene
tables, and the only exchange is the gather of per-configuration results ``north_star`` asks
for -- one ``all_gather_into_tensor`` of ``uint8`` flags (and optionally ``int32`` first-pair
indices) over NCCL/NVLink.  Shards are uniform (the tail is padded), so no gatherv is needed.

The shard / gather logic is backend-agnostic and is exercised on CPU with gloo in
tests/test_distributed.py (the check itself always runs on a GPU).
"""

from __future__ import annotations

import torch
import torch.distributed as dist


def shard_bounds(n: int, world: int, rank: int) -> tuple[int, int, int]:
    """(start, stop, padded shard length) of rank's contiguous shard of n rows."""
    per = (n + world - 1) // world
    start = min(n, rank * per)
    return start, min(n, start + per), per


def gather_shards(local: torch.Tensor, n: int, per: int, fill=0) -> torch.Tensor:
    """All-gather uniform shards ``local[<=per]`` into the full ``[n]`` result on every rank."""
    world = dist.get_world_size() if dist.is_initialized() else 1
    if world == 1:
        return local[:n]
    padded = local
    if local.shape[0] < per:
        pad = torch.full((per - local.shape[0], *local.shape[1:]), fill, dtype=local.dtype, device=local.device)
        padded = torch.cat([local, pad])
    out = torch.empty((world * per, *local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, padded.contiguous())
    return out[:n]


def check_sharded(gpu_model, q_full_or_fn, n: int, *, first_pair: bool = False, check_fn=None) -> dict:
    """Check rank's shard of ``n`` configurations and gather the flags on every rank.

    ``q_full_or_fn``: the full ``[n, ld]`` tensor (each rank slices its shard) or a callable
    ``(start, stop) -> tensor`` that produces only the shard (inputs of large batches are
    generated on the device per rank and never cross PCIe).  ``check_fn`` overrides the
    checker (tests use a CPU stand-in to exercise the plumbing under gloo)."""
    world = dist.get_world_size() if dist.is_initialized() else 1
    rank = dist.get_rank() if dist.is_initialized() else 0
    start, stop, per = shard_bounds(n, world, rank)
    q = q_full_or_fn(start, stop) if callable(q_full_or_fn) else q_full_or_fn[start:stop]
    res = (check_fn or (lambda x: gpu_model.check(x, first_pair=first_pair)))(q)
    out = {"hit": gather_shards(res["hit"].to(torch.uint8), n, per).bool()}
    if first_pair:
        out["first_pair"] = gather_shards(res["first_pair"], n, per, fill=-1)
    return out
