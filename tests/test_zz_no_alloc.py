"""hbt_counts_deltas_static_dev -- the entry point for callers that must not
allocate device memory while sampling (a monitor beside a job that owns almost all
of HBM; a collection loop that cannot take the device-wide stall of cudaMalloc).

Called through the C-ABI twice in a row, on different traces, and compared bit for
bit (integer outputs) with the `deltas` the existing hbt_counts_pipeline_dev
produces for the same inputs, and with the oracle's deltas for them.  The two calls
share the library's fixed scratch array; the second must not be disturbed by what the
first left in it, and it must leave the amount of free device memory exactly as it
found it (the start-up call has resolved the scratch address by then).

This file sorts last on purpose: a run that stops at the first failure has then
already run every other test.  It also passes when run alone in a fresh process.
"""
import errno

import numpy as np
import pytest

import dynolog_1_b200 as hbt
from dynolog_1_b200 import tracegen as tg
import _oracle as orc

pytestmark = pytest.mark.gpu


def reference_deltas(ctx, reads, I, C, E):
    import torch
    out = torch.zeros((I, E), dtype=torch.int64, device="cuda")
    ctx.counts_pipeline_dev(reads, I, C, E, [], deltas=out)
    ctx.sync()
    return out


def test_static_scratch_entry_point_matches_pipeline_twice(built):
    import torch
    torch.cuda.set_device(0)
    stream = torch.cuda.current_stream().cuda_stream
    ctx = hbt.Context(0)
    try:
        # two calls in a row: a large one that fills most of the scratch, then a
        # smaller one with other data whose result must not see the leftovers
        # (3 + E > 32 takes the wide reduce, 3 + E <= 32 the narrow one)
        for call, (seed, (I, C, E)) in enumerate(((11, (8000, 32, 64)), (12, (700, 64, 8)))):
            assert I * E <= hbt.STATIC_SCRATCH_WORDS
            reads = tg.counts_trace(seed, I, C, E, device="cuda", multiplexed=True)
            # one failed read (word0 != n_events, PerCpuBase::read returns false): the
            # interval's scaled counts are 0, so the deltas around it change sign
            reads[I // 3, C // 2, 0] = 0
            got = torch.full((I, E), -1, dtype=torch.int64, device="cuda")
            want = reference_deltas(ctx, reads, I, C, E)   # (allocates: before the measurement)
            host_want = orc.counts(reads.cpu().numpy().view(np.uint64))["deltas"]
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info()[0]
            rc = hbt.counts_deltas_static_dev(reads, I, C, E, got, stream)
            assert rc == 0, rc
            torch.cuda.synchronize()
            free_after = torch.cuda.mem_get_info()[0]
            if call > 0:   # the first call on a device may load the module / resolve the symbol
                assert free_after == free_before, "the allocation-free entry point took device memory"
            assert torch.equal(got, want)  # u64 deltas: bit-exact, no tolerance
            assert np.array_equal(got.cpu().numpy().view(np.uint64), host_want)
            assert int((got != -1).sum()) > 0
    finally:
        ctx.close()


def test_static_scratch_entry_point_refuses_oversize_and_bad_args(built):
    import torch
    torch.cuda.set_device(0)
    E = 8
    I = hbt.STATIC_SCRATCH_WORDS // E + 1  # one interval too many for the scratch
    reads = torch.zeros((1, 1, 3 + E), dtype=torch.int64, device="cuda")
    out = torch.zeros((1, E), dtype=torch.int64, device="cuda")
    assert hbt.counts_deltas_static_dev(reads, I, 1, E, out) == -errno.E2BIG
    assert hbt.counts_deltas_static_dev(reads, 1, 0, E, out) == -errno.EINVAL
    assert hbt.counts_deltas_static_dev(reads, 1, 1, hbt.MAX_EVENTS + 1, out) == -errno.EINVAL
    assert hbt.counts_deltas_static_dev(None, 1, 1, E, out) == -errno.EINVAL
    assert hbt.counts_deltas_static_dev(reads, 0, 1, E, out) == 0  # nothing to do
