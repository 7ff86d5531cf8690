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

from ramp_b200.distributed import ShardedChecker, check_sharded, pack_bits_reference, shard_bounds, unpack_bits


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
    # the streaming form: uniform shards, three batches through two buffer sets, gathers in flight
    per = (n + world - 1) // world
    sc = ShardedChecker(None, per, device=torch.device("cpu"),
                        check_fn=lambda x, hit: hit.copy_(_fake_check(x)["hit"]),
                        pack_fn=lambda hit, words: words.copy_(pack_bits_reference(hit)))
    qp = torch.cat([q, torch.full((world * per - n, 8), -1.0)])  # padding rows are "free"
    tickets, want = [], []
    for k in range(3):
        qk = qp.roll(k, dims=1) * (1.0 if k % 2 == 0 else -1.0)
        tickets.append(sc.submit(qk[rank * per:(rank + 1) * per]))
        want.append(_fake_check(qk)["hit"].bool())
        if k >= 1:  # at most two outstanding: read batch k - 1 while batch k's gather runs
            assert torch.equal(sc.result(tickets[k - 1]), want[k - 1])
    assert torch.equal(sc.result(tickets[2]), want[2])
    sc.drain()
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


def test_bit_packing_layout_round_trip():
    g = torch.Generator().manual_seed(7)
    for n in (1, 31, 32, 33, 1000, 4097):
        hit = (torch.rand(n, generator=g) < 0.3).to(torch.uint8)
        words = pack_bits_reference(hit)
        assert words.dtype == torch.int32 and words.shape[0] == (n + 31) // 32
        assert torch.equal(unpack_bits(words, n), hit.bool())
        r = n - 1  # bit (r & 31) of word r >> 5
        assert bool((int(words[r >> 5]) >> (r & 31)) & 1) == bool(hit[r])


def _nccl_worker(rank, world, port, n, out_dir):
    import numpy as np

    from ramp_b200 import workloads

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    rm, _ = workloads.build("ur5e_1m")
    q = torch.from_numpy(workloads.sample_q(rm.model, n, seed=4242)).cuda()
    res = check_sharded(rm.gpu, q, n, first_pair=True)
    per = (n + world - 1) // world
    sc = ShardedChecker(rm.gpu, per)
    qp = torch.cat([q, q[: world * per - n]])
    t0 = sc.submit(qp[rank * per:(rank + 1) * per].contiguous())
    t1 = sc.submit(qp[rank * per:(rank + 1) * per].contiguous())
    stream = torch.stack([sc.result(t0)[:n], sc.result(t1)[:n]])
    torch.cuda.synchronize()
    torch.save({"hit": res["hit"].cpu(), "first_pair": res["first_pair"].cpu(), "stream": stream.cpu()},
               os.path.join(out_dir, f"r{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_flags_are_byte_identical_to_one_gpu(built, tmp_path, world):
    """SURVEY section 4 test plan (v): the flags and first-pair indices gathered from N real GPUs over
    NCCL equal the single-GPU answer byte for byte (ragged batch: the last shard is short)."""
    if not torch.cuda.is_available() or torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    from ramp_b200 import workloads

    n = 300001
    mp.spawn(_nccl_worker, args=(world, _free_port(), n, str(tmp_path)), nprocs=world, join=True)
    rm, _ = workloads.build("ur5e_1m")
    q = torch.from_numpy(workloads.sample_q(rm.model, n, seed=4242)).cuda()
    ref = rm.gpu.check(q, first_pair=True)
    for r in range(world):
        got = torch.load(tmp_path / f"r{r}.pt")
        assert torch.equal(got["hit"], ref["hit"].cpu())
        assert torch.equal(got["first_pair"], ref["first_pair"].cpu())
        assert torch.equal(got["stream"][0], ref["hit"].cpu()) and torch.equal(got["stream"][1], ref["hit"].cpu())


@pytest.mark.gpu
def test_pack_flags_kernel_matches_the_reference_layout(built):
    from ramp_b200 import workloads

    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    rm, _ = workloads.build("rrr_float")
    g = torch.Generator().manual_seed(11)
    for n in (1, 31, 32, 33, 100000, (1 << 20) + 17):
        hit = (torch.rand(n, generator=g) < 0.4).to(torch.uint8)
        words = torch.full(((n + 31) // 32,), -1, dtype=torch.int32, device="cuda")
        rm.gpu.pack_flags(hit.cuda(), words)
        assert torch.equal(words.cpu(), pack_bits_reference(hit))
