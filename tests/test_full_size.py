"""The CUDA path at BASELINE.json's FULL sizes (256 CPUs x 16 MiB rings; 512 CPUs x 1 M intervals), where the
reference and the C restatement would take minutes: every output is checked against a vectorised
restatement computed with torch on the same device from the generator's own tables (the streams are
regular enough for that: fixed-size samples, LOST records at known places), plus the size-independent
properties the domain offers (sums that telescope, linearity over CPU shards, sortedness of the merged
event stream, idempotence of a second identical batch).  Bit-exact throughout.

The small-size tests compare with the reference itself; this file extends the same claims to the sizes
the benchmark is quoted on."""
import numpy as np
import pytest

import dynolog_1_b200 as hbt
from dynolog_1_b200 import tracegen as tg

pytestmark = pytest.mark.gpu

CPUS, RING = 256, 16 << 20


def _bench():
    import os
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    return bench


def _submit(ctx, imgs, tails, heads):
    bases = np.array([imgs[c].data_ptr() for c in range(imgs.shape[0])], dtype=np.uint64)
    ctx.submit_rings_dev(bases, imgs.shape[1] * 8, np.array(tails, dtype=np.uint64), np.array(heads, dtype=np.uint64))


def test_c3_delta_rows_of_every_cpu(built):
    """Config 3's shape, 8 events: every delta row of every CPU against rec[k] - rec[k-1] computed from
    the generator's table (a LOST in front of a sample re-seeds: no row, one unknown interval)."""
    import torch
    b = _bench()
    dev = torch.device("cuda:0")
    E, lost_every = 8, 5000
    W = 5 + E
    imgs, tails, heads, n_rec = b.make_group_rings(CPUS, RING, 0, dev, E, lost_every=lost_every)
    torch.cuda.synchronize()              # (the library runs on its own stream)
    ctx = hbt.Context(0, max_cpus=CPUS)
    try:
        ctx.ring_configure(CPUS, 0)
        ctx.reset_state(hbt.INVALID_TS)
        _submit(ctx, imgs, tails, heads)
        desc = hbt.sampling_desc(E)
        ctx.decode(desc)
        res = ctx.fetch_results()
        assert ctx.reruns == 0 and ctx.slow_tiles == 0
        assert int(res["n_records"].sum()) == n_rec
        base, off = ctx.rows_packed_dev()
        ctx.sync()                        # (packed on the library's stream)
        total = int(off[CPUS])
        rows = torch.as_tensor(b._DevArray(base, (total, 2 + E), "<i8"), device=dev)
        n = (RING - 65536) // (8 * W + 1)
        k = torch.arange(n, device=dev)
        has_row = (k > 0) & ((k % lost_every) != (lost_every - 1))
        for c in range(CPUS):
            rec = tg.group_sample_stream(tg.stream_seed(3, c), n, E, device=dev).view(n, W)
            want = torch.empty((int(has_row.sum().item()), 2 + E), dtype=torch.int64, device=dev)
            cur, prv = rec[1:][has_row[1:]], rec[:-1][has_row[1:]]
            want[:, 0] = prv[:, 1]
            want[:, 1] = cur[:, 1] - prv[:, 1]
            want[:, 2:] = cur[:, 5:] - prv[:, 5:]
            got = rows[int(off[c]):int(off[c + 1])]
            assert got.shape == want.shape, c
            assert torch.equal(got, want), f"rows of cpu {c} differ"
            assert int(res[c]["num_unknown_count_interval"]) == 1 + n // lost_every, c
            assert int(res[c]["counters"][hbt.CNT_SAMPLE]) == n and int(res[c]["ret"]) == -61   # -ENODATA: drained
            # the rows telescope: durations add up to the time between a run's first and last sample
            assert int(got[:, 1].sum().item()) == int((cur[:, 1] - prv[:, 1]).sum().item())
    finally:
        ctx.close()


def test_c2_histogram_of_every_cpu(built):
    """Config 2 (the headline workload): the (cpu,pid,ip) histogram of every CPU against torch.unique
    over the generator's own keys; a second identical batch doubles every count."""
    import torch
    b = _bench()
    dev = torch.device("cuda:0")
    imgs, tails, heads, n_rec = b.make_c2(CPUS, RING, 0, dev)
    torch.cuda.synchronize()              # (the library runs on its own stream)
    ctx = hbt.Context(0, max_cpus=CPUS)
    try:
        ctx.ring_configure(CPUS, 0)
        ctx.hist_configure((1 << 21) * CPUS)
        desc = hbt.ip_desc(40)
        n = (RING - 65536) // 40
        cdfs = (tg.zipf_cdf(4096, 1.1, dev), tg.zipf_cdf(65536, 1.1, dev))
        for batch in (1, 2):
            ctx.reset_state(hbt.INVALID_TS)
            _submit(ctx, imgs, tails, heads)
            ctx.decode(desc)
            res = ctx.fetch_results()
            assert int(res["n_records"].sum()) == n_rec and ctx.reruns == 0
            ent_p, cur_p, lpc, cap = ctx.hist_lists_dev()
            cursors = torch.as_tensor(b._DevArray(cur_p, (CPUS, lpc), "<u4"), device=dev).to(torch.int64)
            pos = torch.arange(cap, device=dev)[None, :]
            distinct = 0
            for c in range(CPUS):
                s = tg.ip_sample_stream(tg.stream_seed(2, c), n, cpu=c, device=dev, cdfs=cdfs).view(n, 5)
                keys = torch.stack([s[:, 1], s[:, 2] & 0xFFFFFFFF], 1)
                uk, cnt = torch.unique(keys, dim=0, return_counts=True)
                words = torch.as_tensor(b._DevArray(ent_p + c * lpc * cap * 16, (lpc, cap, 2), "<i8"), device=dev)
                live = (pos < cursors[c][:, None])
                e = words[live]                                   # [entries][2]: ip, pid | count << 32
                got = torch.stack([e[:, 0], e[:, 1] & 0xFFFFFFFF, (e[:, 1] >> 32) & 0xFFFFFFFF], 1)
                gk, inv = torch.unique(got[:, :2], dim=0, return_inverse=True)
                assert gk.shape[0] == got.shape[0], f"cpu {c}: a key appears twice after the fold"
                order = torch.argsort(inv)
                assert torch.equal(gk, uk), f"keys of cpu {c} differ"
                assert torch.equal(got[order][:, 2], cnt * batch), f"counts of cpu {c} differ"
                distinct += uk.shape[0]
            assert ctx.hist_size() == distinct
    finally:
        ctx.close()


def test_c4_switch_events_merged_order(built):
    """Config 4's context-switch rings (256 CPUs x 16 MiB, ~115 M records): every CPU's event stream is
    in record order (timestamps never go back), the merged stream is sorted by (tstamp, comp_unit_id) and
    is a permutation of the per-CPU streams (same multiset of timestamps per CPU)."""
    import torch
    b = _bench()
    dev = torch.device("cuda:0")
    imgs, tails, heads, n_rec = b.make_switch_rings(CPUS, RING, 0, dev)
    torch.cuda.synchronize()              # (the library runs on its own stream)
    ctx = hbt.Context(0, max_cpus=CPUS)
    try:
        ctx.ring_configure(CPUS, 0)
        ctx.reset_state(hbt.INVALID_TS)
        _submit(ctx, imgs, tails, heads)
        ctx.decode(hbt.switch_desc(tid_level=1))
        res = ctx.fetch_results()
        assert int(res["n_records"].sum()) == n_rec and ctx.reruns == 0 and ctx.slow_tiles == 0
        base, off = ctx.switch_events_dev()
        ctx.sync()
        total = int(off[CPUS])
        assert total == int(res["n_rows"].sum()) >= n_rec * 0.9
        ev = torch.as_tensor(b._DevArray(base, (total, 2), "<i8"), device=dev)
        ts, cu = ev[:, 0], (ev[:, 1] >> 16) & 0xFFFF
        # per CPU: one comp_unit_id, timestamps in order
        o = torch.as_tensor(off.astype(np.int64), device=dev)
        owner = torch.searchsorted(o, torch.arange(total, device=dev), right=True) - 1
        assert torch.equal(cu, owner)
        same = owner[1:] == owner[:-1]
        per_cpu_in_order = bool((~same | (ts[1:] >= ts[:-1])).all())
        assert per_cpu_in_order
        mbase, n = ctx.switch_events_merged_dev()
        ctx.sync()                        # (merged on the library's stream)
        assert n == total
        m = torch.as_tensor(b._DevArray(mbase, (total, 2), "<i8"), device=dev)
        mts, mcu = m[:, 0], (m[:, 1] >> 16) & 0xFFFF
        in_order = bool(((mts[1:] > mts[:-1]) | ((mts[1:] == mts[:-1]) & (mcu[1:] >= mcu[:-1]))).all())
        assert in_order
        # a permutation: per comp unit the same number of events and the same checksum of both words
        for col in (0, 1):
            a = torch.zeros(CPUS, dtype=torch.int64, device=dev).index_add_(0, cu, ev[:, col])
            c2 = torch.zeros(CPUS, dtype=torch.int64, device=dev).index_add_(0, mcu, m[:, col])
            assert torch.equal(a, c2)
        assert torch.equal(torch.bincount(cu, minlength=CPUS), torch.bincount(mcu, minlength=CPUS))
        # and stable: within a comp unit the merged stream keeps the CPU's own order
        key = mcu * (1 << 40) + torch.arange(total, device=dev)
        back = m[torch.argsort(key)]
        stable = torch.equal(back, ev)
        assert stable
    finally:
        ctx.close()


def test_c5_counter_groups_full_size(built):
    """Config 5's shape (512 CPUs x 1 event x 1 M intervals, 16 GB of reads): sums, scaled counts, rates and
    deltas against torch (u64 adds, IEEE f64 multiply / divide / truncate like getCount), and linearity:
    the sums of two CPU shards add up to the sums of all CPUs (what the multi-GPU all-reduce relies on)."""
    import torch
    dev = torch.device("cuda:0")
    I, Cn, E = 1_000_000, 512, 1
    W = 3 + E
    reads = tg.counts_trace(5, I, Cn, E, device=dev, multiplexed=True)
    ctx = hbt.Context(0, max_cpus=Cn)
    try:
        sums = torch.zeros((I, W), dtype=torch.int64, device=dev)
        scaled = torch.zeros((I, E), dtype=torch.int64, device=dev)
        deltas = torch.zeros((I, E), dtype=torch.int64, device=dev)
        torch.cuda.synchronize()          # (the library runs on its own stream)
        ctx.counts_pipeline_dev(reads, I, Cn, E, [], sums=sums, scaled=scaled, deltas=deltas)
        ctx.sync()
        r = reads.view(I, Cn, W)
        want = r.sum(1)
        want[:, 0] = E      # every CPU read: the marker stays n_events
        assert torch.equal(sums, want)
        en, run, cnt = want[:, 1].double(), want[:, 2].double(), want[:, 3].double()
        ws = torch.where(want[:, 1] == 0, torch.zeros_like(en), cnt * en / run).to(torch.int64)
        assert torch.equal(scaled[:, 0], ws)
        wd = ws.clone()
        wd[1:] = ws[1:] - ws[:-1]
        assert torch.equal(deltas[1:, 0], wd[1:])
        assert int(deltas[1:, 0].sum().item()) == int((ws[-1] - ws[0]).item())   # the deltas telescope
        # linearity over CPU shards
        half = Cn // 2
        s0 = torch.zeros_like(sums)
        s1 = torch.zeros_like(sums)
        for shard, out in ((slice(0, half), s0), (slice(half, Cn), s1)):
            part = reads.view(I, Cn, W)[:, shard].contiguous()
            torch.cuda.synchronize()      # (the library runs on its own stream)
            ctx.counts_reduce_dev(part, I, half, E, out)
            ctx.sync()
            del part
        tot = s0 + s1
        tot[:, 0] = E
        assert torch.equal(tot, sums)
    finally:
        ctx.close()


@pytest.mark.parametrize("switch", [False, True], ids=["mixed_rows", "switch_events"])
def test_c4_both_kernels_agree(built, switch):
    """Config 4's streams at full size through the two independent decode kernels (k_decode: a CTA per
    segment, tiles, guesses checked across sub-chunks; k_walk: a warp per segment, lanes following
    header.size): the same per-CPU results, the same delta rows / events, bit for bit.  (At small sizes both
    are held to the reference; its consume() needs minutes for 4 GiB.)"""
    import torch
    b = _bench()
    dev = torch.device("cuda:0")
    E = 8
    if switch:
        imgs, tails, heads, n_rec = b.make_switch_rings(CPUS, RING, 0, dev)
        desc = hbt.switch_desc(tid_level=1)
    else:
        imgs, tails, heads, n_rec = b.make_mixed_rings(CPUS, RING, 0, dev, E)
        desc = hbt.sampling_desc(E, handled=hbt.MODE_MIXED)
    torch.cuda.synchronize()
    out = []
    for kernel in (hbt.KERNEL_TILES, hbt.KERNEL_WALK):
        ctx = hbt.Context(0, max_cpus=CPUS)
        try:
            ctx.set_decode_kernel(kernel)
            ctx.ring_configure(CPUS, 0)
            ctx.reset_state(hbt.INVALID_TS)
            _submit(ctx, imgs, tails, heads)
            ctx.decode(desc)
            res = ctx.fetch_results().copy()
            assert int(res["n_records"].sum()) == n_rec and ctx.reruns == 0
            if switch:
                base, off = ctx.switch_events_dev()
                width = 2
            else:
                base, off = ctx.rows_packed_dev()
                width = 2 + E
            ctx.sync()
            data = torch.as_tensor(b._DevArray(base, (int(off[CPUS]), width), "<i8"), device=dev).clone()
            out.append((res, off.copy(), data))
        finally:
            ctx.close()
    (r0, o0, d0), (r1, o1, d1) = out
    assert r0.tobytes() == r1.tobytes()
    assert np.array_equal(o0, o1)
    same = torch.equal(d0, d1)
    assert same
    assert d0.shape[0] > n_rec * 0.5
