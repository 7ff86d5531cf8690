"""World-size-2 gloo test of the shard / gather plumbing (runs on CPU).

The collision check itself needs a GPU, so a deterministic CPU stand-in plays the checker; what
is tested is that shards tile the batch exactly, ragged tails are padded and trimmed, and the
gathered flags are identical on every rank and equal to the single-process result (shard
invariance)."""

import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ramp_b200.distributed import check_sharded, shard_bounds


def _fake_check(q):
    hit = (q.sum(dim=1) > 0).to(torch.uint8)
    first = torch.where(hit.bool(), (q[:, 0].abs() * 10).to(torch.int32) % 7, torch.full_like(hit, -1, dtype=torch.int32))
    return {"hit": hit, "first_pair": first}


def _worker(rank, world, port, n, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    g = torch.Generator().manual_seed(1234)
    q = torch.randn((n, 8), generator=g)
    res = check_sharded(None, q, n, first_pair=True, check_fn=_fake_check)
    res2 = check_sharded(None, lambda a, b: q[a:b], n, check_fn=_fake_check)
    assert torch.equal(res["hit"], res2["hit"])
    torch.save(res, os.path.join(out_dir, f"r{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("n", [1000, 1001, 1, 2])
def test_shard_gather_world2(tmp_path, n):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), n, str(tmp_path)), nprocs=world, join=True)
    g = torch.Generator().manual_seed(1234)
    ref = _fake_check(torch.randn((n, 8), generator=g))
    for r in range(world):
        got = torch.load(tmp_path / f"r{r}.pt")
        assert torch.equal(got["hit"], ref["hit"].bool())
        assert torch.equal(got["first_pair"], ref["first_pair"])


def test_shard_bounds_tile_the_batch():
    for n in (0, 1, 5, 31, 32, 33, 1000, 1 << 20):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_bounds(n, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            for (a0, a1, per), (b0, _, _) in zip(spans, spans[1:]):
                assert a1 == b0 and a1 - a0 <= per
            assert sum(b - a for a, b, _ in spans) == n
