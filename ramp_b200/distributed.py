"""Batch sharding across the GPUs of one box.

Configurations are independent units: rank r checks the contiguous shard
``[r * ceil(N / R), ...)`` on its own GPU with its own replica of the (tiny) robot / scene
tables, and the only exchange is the gather of per-configuration results ``north_star`` asks
for.  The flags travel as ONE BIT per configuration (``ramp_pack_flags``: a ballot per 32
flags), all-gathered over NCCL / NVLink; first-pair indices, when asked for, as ``int32``.
Shards are uniform (the tail is padded), so no gatherv is needed.

Two entry points: ``check_sharded`` (one call, blocking: check, gather, return the full result
on every rank) and ``ShardedChecker`` (a stream of batches: the gather of batch i is issued
asynchronously and runs while batch i + 1 is being checked; two result buffers).

The shard / gather logic is backend-agnostic and is exercised on CPU with gloo in
tests/test_distributed.py (the check itself always runs on a GPU; ``check_fn`` / ``pack_fn``
let the test stand in for the device calls).
"""

from __future__ import annotations

import torch
import torch.distributed as dist


def _world() -> tuple[int, int]:
    if dist.is_available() and dist.is_initialized():
        return dist.get_world_size(), dist.get_rank()
    return 1, 0


def shard_bounds(n: int, world: int, rank: int) -> tuple[int, int, int]:
    """(start, stop, padded shard length) of rank's contiguous shard of n rows."""
    per = (n + world - 1) // world
    start = min(n, rank * per)
    return start, min(n, start + per), per


def pack_bits_reference(hit: torch.Tensor) -> torch.Tensor:
    """The layout ``ramp_pack_flags`` produces, in torch: bit (r & 31) of word r >> 5.  Host-side
    helper for CPU tensors (tests, unpacking bookkeeping) -- the device path is the CUDA kernel."""
    n = hit.shape[0]
    pad = (-n) % 32
    b = torch.cat([hit.to(torch.int64) != 0, torch.zeros(pad, dtype=torch.bool, device=hit.device)]).view(-1, 32)
    w = (b.to(torch.int64) << torch.arange(32, device=hit.device, dtype=torch.int64)).sum(dim=1)
    return torch.where(w >= 2 ** 31, w - 2 ** 32, w).to(torch.int32)


def unpack_bits(words: torch.Tensor, n: int) -> torch.Tensor:
    """int32 bit words -> bool[n]."""
    bits = (words.view(-1, 1) >> torch.arange(32, device=words.device, dtype=torch.int32)) & 1
    return bits.view(-1)[:n].bool()


def gather_shards(local: torch.Tensor, n: int, per: int, fill=0) -> torch.Tensor:
    """All-gather uniform shards ``local[<=per]`` into the full ``[n]`` result on every rank."""
    world, _ = _world()
    if world == 1:
        return local[:n]
    padded = local
    if local.shape[0] < per:
        pad = torch.full((per - local.shape[0], *local.shape[1:]), fill, dtype=local.dtype, device=local.device)
        padded = torch.cat([local, pad])
    out = torch.empty((world * per, *local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, padded.contiguous())
    return out[:n]


class ShardedChecker:
    """Checks this rank's shard of successive batches and gathers the flags of all ranks.

    Every rank passes shards of the same length ``n_local`` (pad the last rank's).  ``submit(q)``
    checks the shard (``ramp_check_batch``), packs the flags to bits (``ramp_pack_flags``) and
    issues the all-gather asynchronously; it returns a ticket.  ``result(ticket)`` waits for that
    gather and returns ``bool[world * n_local]`` (rank-major, i.e. global row order).  Two buffer
    sets alternate, so at most two tickets are outstanding: the gather of one batch runs while
    the next one is checked."""

    def __init__(self, gpu_model, n_local: int, *, device=None, check_fn=None, pack_fn=None):
        self.gpu = gpu_model
        self.n_local = int(n_local)
        self.world, self.rank = _world()
        dev = device if device is not None else (gpu_model.device if gpu_model is not None else torch.device("cpu"))
        self.words = (self.n_local + 31) // 32
        self.hits = [torch.empty(self.n_local, dtype=torch.uint8, device=dev) for _ in range(2)]
        self.packed = [torch.zeros(self.words, dtype=torch.int32, device=dev) for _ in range(2)]
        self.gathered = [torch.empty(self.words * self.world, dtype=torch.int32, device=dev) for _ in range(2)]
        self.pending = [None, None]
        self.turn = 0
        self._check = check_fn or (lambda q, hit: self.gpu.check_raw(q, hit))
        self._pack = pack_fn or (lambda hit, words: self.gpu.pack_flags(hit, words))

    def submit(self, q_local: torch.Tensor) -> int:
        assert q_local.shape[0] == self.n_local, "every rank submits shards of n_local rows"
        b = self.turn
        self.turn ^= 1
        if self.pending[b] is not None:  # the gather that read this buffer set two submits ago
            self.pending[b].wait()
            self.pending[b] = None
        self._check(q_local, self.hits[b])
        self._pack(self.hits[b], self.packed[b])
        if self.world > 1:
            self.pending[b] = dist.all_gather_into_tensor(self.gathered[b], self.packed[b], async_op=True)
        return b

    def wait(self, ticket: int) -> None:
        if self.pending[ticket] is not None:
            self.pending[ticket].wait()
            self.pending[ticket] = None

    def drain(self) -> None:
        for b in range(2):
            self.wait(b)

    def local_flags(self, ticket: int) -> torch.Tensor:
        return self.hits[ticket]

    def result(self, ticket: int) -> torch.Tensor:
        """bool[world * n_local]: the flags of every rank's shard, in rank order."""
        self.wait(ticket)
        words = self.gathered[ticket] if self.world > 1 else self.packed[ticket]
        per_rank = unpack_bits(words, self.words * 32 * self.world).view(self.world, self.words * 32)
        return per_rank[:, : self.n_local].reshape(-1)


def check_sharded(gpu_model, q_full_or_fn, n: int, *, first_pair: bool = False, check_fn=None) -> dict:
    """Check rank's shard of ``n`` configurations and gather the results on every rank.

    ``q_full_or_fn``: the full ``[n, ld]`` tensor (each rank slices its shard) or a callable
    ``(start, stop) -> tensor`` that produces only the shard (inputs of large batches are
    generated on the device per rank and never cross PCIe).  ``check_fn`` overrides the
    checker (tests use a CPU stand-in to exercise the plumbing under gloo).  Flags are gathered
    as bits, first-pair indices as int32."""
    world, rank = _world()
    start, stop, per = shard_bounds(n, world, rank)
    q = q_full_or_fn(start, stop) if callable(q_full_or_fn) else q_full_or_fn[start:stop]
    res = (check_fn or (lambda x: gpu_model.check(x, first_pair=first_pair)))(q)
    hit = res["hit"].to(torch.uint8).contiguous()
    words_per = (per + 31) // 32
    packed = torch.zeros(words_per, dtype=torch.int32, device=hit.device)
    if hit.shape[0]:
        if hit.is_cuda and gpu_model is not None:
            gpu_model.pack_flags(hit, packed)
        else:
            packed[: (hit.shape[0] + 31) // 32] = pack_bits_reference(hit)
    if world > 1:
        allw = torch.empty(words_per * world, dtype=torch.int32, device=hit.device)
        dist.all_gather_into_tensor(allw, packed)
    else:
        allw = packed
    flags = unpack_bits(allw, words_per * 32 * world).view(world, words_per * 32)[:, :per].reshape(-1)[:n]
    out = {"hit": flags}
    if first_pair:
        out["first_pair"] = gather_shards(res["first_pair"], n, per, fill=-1)
    return out
