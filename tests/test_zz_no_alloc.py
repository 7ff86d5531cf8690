"""ramp_check_batch_noalloc: the allocation-free variant of the collision check.

It serves hosts that may not allocate device memory while they run -- a planner thread in a
real-time loop (an allocation can stall every stream of the process for an unknown time) or a
process that handed out all of its device memory at start-up.  The entry point must give
exactly what ramp_check_batch gives; its work queue is a fixed array inside the library that
survives from call to call, so the check is done twice in a row with different data to show
that nothing of the first call leaks into the second.  The file name sorts last on purpose:
with `pytest -x` every other test has already run when this one does.
"""

import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

MAX_ROWS = 32768  # RAMP_NOALLOC_MAX_ROWS in include/ramp_b200.h


def _noalloc(gpu, q, *, first_pair=False, pair_mask=False):
    """Call the entry point through the C ABI with caller-owned tensors only."""
    n = q.shape[0]
    npairs = gpu.info()["n_pairs"]
    hit = torch.full((n,), 7, dtype=torch.uint8, device=q.device)
    first = torch.full((n,), 12345, dtype=torch.int32, device=q.device) if first_pair else None
    words = max((npairs + 31) // 32, 1)
    mask = torch.zeros((n, words), dtype=torch.int32, device=q.device) if pair_mask else None
    stream = torch.cuda.current_stream(q.device).cuda_stream
    rc = gpu.lib.ramp_check_batch_noalloc(gpu.handle, q.data_ptr(), n, q.shape[1], 2 if pair_mask else 0, hit.data_ptr(),
                                          None if first is None else first.data_ptr(),
                                          None if mask is None else mask.data_ptr(), stream)
    return rc, hit, first, mask


def _unpack(mask, npairs):
    bits = (mask[:, :, None] >> torch.arange(32, device=mask.device, dtype=torch.int32)) & 1
    return bits.reshape(mask.shape[0], -1)[:, :npairs].bool()


@pytest.mark.parametrize("workload", ["ur5e_1m", "panda_4096"])
def test_noalloc_matches_check_batch_twice_in_a_row(built, workload):
    from ramp_b200 import workloads

    rm, _ = workloads.build(workload)
    gpu = rm.gpu
    npairs = gpu.info()["n_pairs"]
    torch.cuda.synchronize()
    # two back-to-back calls on one stream, different inputs, the second one smaller, so stale
    # queue entries of the first call lie beyond the second call's counter
    for seed, n in ((21, MAX_ROWS), (22, 20000)):
        q = torch.from_numpy(workloads.sample_q(rm.model, n, seed=seed)).cuda()
        ref = gpu.check(q, first_pair=True, pair_mask=True)
        ref_fast = gpu.check(q, first_pair=True)
        rc, hit, first, mask = _noalloc(gpu, q, first_pair=True, pair_mask=True)
        assert rc == 0, gpu.lib.ramp_last_error()
        rc2, hit2, first2, _ = _noalloc(gpu, q, first_pair=True)  # early-out mode
        assert rc2 == 0, gpu.lib.ramp_last_error()
        torch.cuda.synchronize()
        # integer / boolean outputs: bit exact, the tolerance the parity tests use for them
        assert torch.equal(hit.bool(), ref["hit"])
        assert torch.equal(first, ref["first_pair"])
        assert torch.equal(_unpack(mask, npairs), ref["pair_mask"])
        assert torch.equal(hit2.bool(), ref_fast["hit"])
        assert torch.equal(first2, ref_fast["first_pair"])
        assert ref["hit"].any() and not ref["hit"].all()  # the comparison is not vacuous


def test_noalloc_rejects_oversize_and_reports_why(built):
    from ramp_b200 import workloads

    rm, _ = workloads.build("panda_4096")
    gpu = rm.gpu
    q = torch.zeros((MAX_ROWS + 1, workloads.padded_nq(rm.model)), dtype=torch.float32, device="cuda")
    rc, hit, _, _ = _noalloc(gpu, q)
    assert rc == -2  # RAMP_ERR_LIMIT
    msg = gpu.lib.ramp_last_error().decode()
    assert "ramp_check_batch_noalloc" in msg and str(MAX_ROWS) in msg
    torch.cuda.synchronize()
    assert (hit == 7).all()  # nothing was launched
    # empty batch: accepted, nothing to do
    rc, _, _, _ = _noalloc(gpu, q[:0])
    assert rc == 0


def test_noalloc_primitive_robot(built):
    """A robot without GJK pairs (boxes, cylinder, sphere) goes through the same entry point."""
    from ramp_b200 import workloads

    rm, _ = workloads.build("rrr_float")
    gpu = rm.gpu
    q = torch.from_numpy(workloads.sample_q(rm.model, 5000, seed=5)).cuda()
    ref = gpu.check(q, first_pair=True)
    rc, hit, first, _ = _noalloc(gpu, q, first_pair=True)
    assert rc == 0
    torch.cuda.synchronize()
    assert torch.equal(hit.bool(), ref["hit"]) and torch.equal(first, ref["first_pair"])
