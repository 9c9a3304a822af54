"""Parity of the CUDA path (through the C-ABI, libhbt_b200.so) with the oracle on the same
seeded inputs.  orc.consume() runs the REFERENCE's own consume() / generators (oracle/_ref, which
travels to the GPU box) wherever they cover the decode and cross-checks the C restatement
(oracle/hbt_oracle.c) against them; without oracle/_ref the restatement (pinned by
test_oracle_vs_reference.py / test_golden.py) is the checker.
Bit-exact for every integer output; f64 ratios within 1e-12 relative."""
import numpy as np
import pytest

import dynolog_1_b200 as hbt
from dynolog_1_b200 import tracegen as tg
import _oracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", params=[hbt.KERNEL_TILES, hbt.KERNEL_WALK], ids=["tiles", "walk"])
def ctx(built, request):
    """Every test runs under both decode kernels (the choice only matters where k_walk covers the
    aggregation; results must never depend on it)."""
    c = hbt.Context(0)
    c.set_decode_kernel(request.param)
    yield c
    c.close()


def ring_np(stream, ds, tail, fill=77):
    ring, head = tg.place_in_ring(stream, ds, tail, fill_seed=fill)
    return ring.numpy().view(np.uint64), head


def gpu_decode(ctx, rings, desc, ds, seed_ts=hbt.INVALID_TS, reset=True, hist_slots=0):
    if reset:
        ctx.ring_configure(len(rings), ds)
        ctx.reset_state(seed_ts)
        if hist_slots:
            ctx.hist_configure(hist_slots)
    for cpu, (ring, tail, head) in enumerate(rings):
        ctx.submit_ring_image(cpu, ring, tail, head)
    ctx.decode(desc)
    return ctx.fetch_results()


def gpu_decode_full_images(ctx, rings, desc, ds, seed_ts=hbt.INVALID_TS, reset=True):
    """Whole ring images on the device (hbt_submit_ring_dev): bytes outside
    [tail, head) are there too, like in the kernel's mmap, so a record that starts
    before head and runs past it reads the same bytes as in the reference."""
    import torch
    imgs = [torch.from_numpy(r.view(np.int64)).cuda() for r, _, _ in rings]
    if reset:
        ctx.ring_configure(len(rings), 0)
        ctx.reset_state(seed_ts)
    for cpu, (ring, tail, head) in enumerate(rings):
        ctx.submit_ring_dev(cpu, imgs[cpu], ds, tail, head)
    ctx.decode(desc)
    res = ctx.fetch_results()
    ctx._keep = imgs
    return res


def check_cpu(ctx, cpu, got, want, desc):
    assert int(got["ret"]) == want["ret"], (cpu, int(got["ret"]), want["ret"])
    assert int(got["n_records"]) == want["n_records"], cpu
    assert int(got["new_tail"]) == want["new_tail"], cpu
    assert [int(x) for x in got["counters"]] == want["counters"], cpu
    if desc.aggregate == hbt.AGG_COUNT_DELTAS:
        assert int(got["n_rows"]) == want["n_rows"], cpu
        assert int(got["num_unknown_count_interval"]) == want["num_unknown"], cpu
        assert int(got["carry_tstamp"]) == want["carry_tstamp"], cpu
        rows = ctx.fetch_rows(cpu, int(got["n_rows"]), desc.n_events)
        assert np.array_equal(rows, want["rows"]), f"rows differ on cpu {cpu}"
        if want["counters"][hbt.CNT_SAMPLE]:
            assert np.array_equal(ctx.fetch_carry(cpu, desc.n_events), want["carry_last"]), cpu


# ---------------------------------------------------------------------------
# counter groups (K4 + K5)
# ---------------------------------------------------------------------------
@pytest.mark.parametrize("I,C,E,mux", [(300, 64, 8, False), (200, 64, 8, True), (64, 7, 32, True),
                                       (100, 3, 1, True), (9, 512, 2, True), (40, 16, 64, True),
                                       (1, 1, 1, False), (1000, 256, 32, True)])
def test_counts_pipeline(ctx, I, C, E, mux):
    reads = tg.counts_trace(61 + E, I, C, E, multiplexed=mux).numpy().view(np.uint64)
    ratios = [(min(1, E - 1), 0), (E - 1, min(1, E - 1))]
    got, want = ctx.counts_pipeline_host(reads, ratios), orc.counts(reads, ratios)
    for k in ("sums", "scaled", "deltas"):
        assert np.array_equal(got[k], want[k]), k
    assert np.array_equal(got["rates"], want["rates"])
    # tolerance stated by the north star for f64 ratios
    assert np.allclose(got["ratios"], want["ratios"], rtol=1e-12, atol=0.0)
    assert np.array_equal(got["ratios"], want["ratios"])  # and in fact identical (IEEE div)


def test_counts_edge_values(ctx):
    I, C, E = 6, 4, 3
    reads = tg.counts_trace(71, I, C, E, multiplexed=True).numpy().view(np.uint64).copy()
    reads[1, :, 3] = np.uint64(2**62)
    reads[2, :, 4] = np.uint64(2**60 + 12345)
    reads[3, :, 1] = 0
    reads[3, :, 2] = 0
    reads[4, 2, 0] = 0
    got, want = ctx.counts_pipeline_host(reads), orc.counts(reads)
    assert np.array_equal(got["sums"], want["sums"])
    assert got["sums"][4, 0] == 0
    assert np.array_equal(got["scaled"], want["scaled"])
    assert np.array_equal(got["deltas"], want["deltas"])
    ok = [0, 1, 2, 5]  # rate of a 0/0 interval is NaN on both sides
    assert np.array_equal(got["rates"][ok], want["rates"][ok])
    assert np.all(np.isnan(got["rates"][3])) and np.all(np.isnan(want["rates"][3]))


def test_counts_dev_split_equals_pipeline(ctx):
    """K4 then K5 (the multi-GPU order: reduce, [NCCL sum], scale) == fused."""
    import torch
    I, C, E = 500, 128, 8
    reads = tg.counts_trace(5, I, C, E, device="cuda", multiplexed=True)
    sums = torch.zeros((I, 3 + E), dtype=torch.int64, device="cuda")
    scaled = torch.zeros((I, E), dtype=torch.int64, device="cuda")
    rates = torch.zeros((I, E), dtype=torch.float32, device="cuda")
    ratio = torch.zeros((I, 1), dtype=torch.float64, device="cuda")
    deltas = torch.zeros((I, E), dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    prev_stream = torch.cuda.current_stream()
    torch.cuda.set_stream(torch.cuda.Stream())
    ctx.use_torch_stream()
    try:
        ctx.counts_reduce_dev(reads, I, C, E, sums)
        ctx.counts_scale_dev(sums, I, E, [(1, 0)], None, scaled, rates, ratio, deltas)
        ctx.sync()
        want = orc.counts(reads.cpu().numpy().view(np.uint64), [(1, 0)])
        assert np.array_equal(sums.cpu().numpy().view(np.uint64), want["sums"])
        assert np.array_equal(scaled.cpu().numpy().view(np.uint64), want["scaled"])
        assert np.array_equal(deltas.cpu().numpy().view(np.uint64), want["deltas"])
        assert np.array_equal(rates.cpu().numpy(), want["rates"])
        assert np.allclose(ratio.cpu().numpy(), want["ratios"], rtol=1e-12, atol=0.0)
        # linearity: sums of a trace split by CPU add up (what the NCCL reduce relies on)
        s0 = torch.zeros_like(sums)
        s1 = torch.zeros_like(sums)
        ctx.counts_reduce_dev(reads[:, :C // 2].contiguous(), I, C // 2, E, s0)
        ctx.counts_reduce_dev(reads[:, C // 2:].contiguous(), I, C // 2, E, s1)
        ctx.sync()
        tot = s0 + s1
        tot[:, 0] = E
        assert torch.equal(tot, sums)
    finally:
        ctx.set_stream(None)
        torch.cuda.synchronize()
        torch.cuda.set_stream(prev_stream)


# ---------------------------------------------------------------------------
# ring decode + per-interval deltas (K1..K3)
# ---------------------------------------------------------------------------
@pytest.mark.parametrize("n_events,lost_every,ds,n_cpus", [
    (8, 0, 1 << 17, 3), (8, 50, 1 << 17, 3), (1, 7, 1 << 16, 2), (32, 13, 1 << 18, 2),
    (64, 0, 1 << 18, 2), (8, 1000, 1 << 22, 5), (32, 400, 1 << 23, 2), (2, 3, 1 << 20, 4)])
def test_sample_deltas(ctx, n_events, lost_every, ds, n_cpus):
    desc = hbt.sampling_desc(n_events)
    rings = []
    for cpu in range(n_cpus):
        n = (ds - 4096 - 512 * cpu) // (40 + 8 * n_events + 24)
        s = tg.group_sample_stream(tg.stream_seed(3, cpu) + n_events, n, n_events,
                                   lost_every=lost_every)
        tail = tg.random_tail(n_events + lost_every + cpu, ds)
        ring, head = ring_np(s, ds, tail)
        rings.append((ring, tail, head))
    res = gpu_decode(ctx, rings, desc, ds)
    wants = [orc.consume(ring, tail, head, desc) for ring, tail, head in rings]
    for cpu in range(n_cpus):
        check_cpu(ctx, cpu, res[cpu], wants[cpu], desc)
    assert ctx.reruns == 0 and ctx.slow_tiles == 0, "well-formed streams must stay on the fast path"
    check_rows_all(ctx, wants, n_events)


def check_rows_all(ctx, wants, n_events):
    """hbt_fetch_rows_all: every CPU's rows packed on the device, one copy for the whole decode."""
    rows, off = ctx.fetch_rows_all(n_events)
    assert int(off[-1]) == sum(w["n_rows"] for w in wants) == len(rows)
    for cpu, w in enumerate(wants):
        assert np.array_equal(rows[int(off[cpu]):int(off[cpu + 1])], w["rows"]), f"packed rows of cpu {cpu}"


def test_carry_across_calls(ctx):
    """A window decoded in three calls (new data submitted between them) equals
    the oracle chained the same way: counters, carry tstamp / last_[] and rows."""
    ds, E = 1 << 18, 4
    s = tg.group_sample_stream(11, 3000, E, lost_every=170)
    tail0 = tg.random_tail(5, ds)
    ring, head = ring_np(s, ds, tail0)
    cuts = [tail0, tail0 + 8 * (5 + E) * 1000, tail0 + 8 * (5 + E) * 1001 + 24 * 5, head]
    desc = hbt.sampling_desc(E)
    ctx.ring_configure(1, 0)
    ctx.reset_state(777)
    st = dict(carry_tstamp=777, carry_last=np.zeros(E, np.uint64), counters=[0] * 11, num_unknown=0)
    a = tail0
    for b in cuts[1:]:
        # heads in the middle of a record: consume() takes the straddling record
        # and the next window starts where it left the tail
        if b <= a:
            continue
        want = orc.consume(ring, a, b, desc, st["carry_tstamp"], st["carry_last"],
                                st["counters"], st["num_unknown"])
        got = gpu_decode_full_images(ctx, [(ring, a, b)], desc, ds, reset=False)[0]
        check_cpu(ctx, 0, got, want, desc)
        st = dict(carry_tstamp=want["carry_tstamp"], carry_last=want["carry_last"],
                  counters=want["counters"], num_unknown=want["num_unknown"])
        a = want["new_tail"]


@pytest.mark.parametrize("max_records", [0, 1, 100, 4321, 10**6])
def test_max_records(ctx, max_records):
    ds, E = 1 << 21, 8
    rings = []
    for cpu in range(3):
        s = tg.group_sample_stream(100 + cpu, 15000 + 100 * cpu, E, lost_every=333)
        tail = tg.random_tail(cpu + 40, ds)
        ring, head = ring_np(s, ds, tail)
        rings.append((ring, tail, head))
    desc = hbt.sampling_desc(E, max_records=max_records)
    res = gpu_decode(ctx, rings, desc, ds)
    wants = [orc.consume(ring, tail, head, desc) for ring, tail, head in rings]
    for cpu in range(len(rings)):
        check_cpu(ctx, cpu, res[cpu], wants[cpu], desc)
    check_rows_all(ctx, wants, E)   # rows of a window cut by max_num_records


def test_timestamp_regression_and_bad_size(ctx):
    ds, E = 1 << 20, 8
    desc = hbt.sampling_desc(E)
    cases = []
    for at in (0, 1, 157, 158, 5000, 9998):  # in-tile, at tile and segment seams
        s = tg.group_sample_stream(21 + at, 9999, E, regress_at=at if at else -1)
        if at == 0:
            s[(5 + E) * 4000 + 2] = E + 1  # header.size != 40 + 8*nr
        cases.append(s)
    rings = []
    for i, s in enumerate(cases):
        tail = tg.random_tail(i + 3, ds)
        ring, head = ring_np(s, ds, tail)
        rings.append((ring, tail, head))
    res = gpu_decode(ctx, rings, desc, ds)
    for cpu, (ring, tail, head) in enumerate(rings):
        want = orc.consume(ring, tail, head, desc)
        assert want["ret"] == -1
        check_cpu(ctx, cpu, res[cpu], want, desc)


def test_regression_at_every_position_small(ctx):
    """Regression swept over every sample of a stream that spans several tiles and
    two segments' worth of seams."""
    ds, E = 1 << 16, 1
    desc = hbt.sampling_desc(E)
    for at in list(range(1, 12)) + [340, 341, 342, 343, 683, 1000]:
        s = tg.group_sample_stream(900 + at, 1200, E, regress_at=at)
        tail = 48 * 7
        ring, head = ring_np(s, ds, tail)
        res = gpu_decode(ctx, [(ring, tail, head)], desc, ds)
        check_cpu(ctx, 0, res[0], orc.consume(ring, tail, head, desc), desc)


def test_tsc_transform(ctx):
    ds, E = 1 << 19, 6
    s = tg.group_sample_stream(31, 4000, E, lost_every=90)
    tail = 4096 * 3 + 64
    ring, head = ring_np(s, ds, tail)
    desc = hbt.sampling_desc(E, tsc_mask=0b100101, tsc=(987654321, 1234567, 21))
    res = gpu_decode(ctx, [(ring, tail, head)], desc, ds)
    check_cpu(ctx, 0, res[0], orc.consume(ring, tail, head, desc), desc)


@pytest.mark.parametrize("mode", [hbt.MODE_SAMPLING, hbt.MODE_CONTEXT_SWITCH, hbt.MODE_AUX_SPACE,
                                  hbt.MODE_MIXED])
@pytest.mark.parametrize("max_records", [hbt.DRAIN, 7777])
@pytest.mark.parametrize("mix", [tg.MIXED_MIX, tg.FULL_MIX], ids=["config4", "all_types"])
def test_mixed_stream_counters(ctx, mode, max_records, mix):
    """Every record type consume() dispatches on (the all_types mix adds THROTTLE / UNTHROTTLE /
    READ / AUX / ITRACE_START, CpuEventsGroup.h:1409-1420, 1446-1496) under every handler set."""
    ds, E = 1 << 21, 8
    rings = []
    for cpu in range(4):
        s = tg.mixed_stream(41 + cpu, 18000, E, mix=mix)
        tail = tg.random_tail(9 + cpu, ds)
        ring, head = ring_np(s, ds, tail)
        rings.append((ring, tail, head))
    desc = hbt.count_desc(mode, max_records=max_records)
    res = gpu_decode(ctx, rings, desc, ds)
    for cpu, (ring, tail, head) in enumerate(rings):
        check_cpu(ctx, cpu, res[cpu], orc.consume(ring, tail, head, desc), desc)


def first_record_of_type(s, rtype, nth):
    """Word index of the nth record of PERF type `rtype` in a stream of u64 words."""
    w = s.numpy().view(np.uint64)
    pos, k = 0, 0
    while pos < len(w):
        h = int(w[pos])
        if (h & 0xFFFFFFFF) == rtype:
            k += 1
            if k == nth:
                return pos
        pos += (h >> 48) // 8
    raise AssertionError("not enough records of that type")


@pytest.mark.parametrize("mode", [hbt.MODE_SAMPLING, hbt.MODE_MIXED, hbt.MODE_CONTEXT_SWITCH])
def test_read_record_with_bad_size(ctx, mode):
    """A PERF_RECORD_READ whose header.size disagrees with its nr: -EPERM in front of it where
    READ has a handler (CpuEventsGroup.h:1446-1456), skipped by size where it has none."""
    ds, E = 1 << 20, 8
    rings = []
    for cpu, nth in enumerate((1, 40, 333)):     # first tile, a later tile, another segment
        s = tg.mixed_stream(141 + cpu, 9000, E, mix=tg.FULL_MIX).clone()
        s[first_record_of_type(s, tg.REC_READ, nth) + 2] = E + 3
        tail = tg.random_tail(19 + cpu, ds)
        ring, head = ring_np(s, ds, tail)
        rings.append((ring, tail, head))
    desc = hbt.count_desc(mode)
    res = gpu_decode(ctx, rings, desc, ds)
    for cpu, (ring, tail, head) in enumerate(rings):
        want = orc.consume(ring, tail, head, desc)
        assert (want["ret"] == -1) == bool(mode & hbt.H_READ)
        check_cpu(ctx, cpu, res[cpu], want, desc)


def test_resume_from_saved_state(built):
    """Checkpoint / resume: the state a decode leaves (counters, count_.tstamp, last_[],
    num_unknown) read out of one context and restored into a fresh one with hbt_ring_set_carry;
    the second half decoded there equals the oracle chained across the cut."""
    ds, E = 1 << 19, 6
    s = tg.group_sample_stream(77, 5000, E, lost_every=211)
    tail = tg.random_tail(6, ds)
    ring, head = ring_np(s, ds, tail)
    desc1 = hbt.sampling_desc(E, max_records=2600)
    desc2 = hbt.sampling_desc(E)
    a, b = hbt.Context(0), hbt.Context(0)
    try:
        r1 = gpu_decode(a, [(ring, tail, head)], desc1, ds, seed_ts=4242)[0]
        w1 = orc.consume(ring, tail, head, desc1, carry_tstamp=4242, carry_last=np.zeros(E, np.uint64))
        check_cpu(a, 0, r1, w1, desc1)
        last = a.fetch_carry(0, E)
        b.ring_configure(1, ds)
        b.set_carry(0, int(r1["carry_tstamp"]), last, r1["counters"], int(r1["num_unknown_count_interval"]))
        b.submit_ring_image(0, ring, int(r1["new_tail"]), head)
        b.decode(desc2)
        r2 = b.fetch_results()[0]
        w2 = orc.consume(ring, w1["new_tail"], head, desc2, w1["carry_tstamp"], w1["carry_last"],
                         w1["counters"], w1["num_unknown"])
        check_cpu(b, 0, r2, w2, desc2)
    finally:
        a.close()
        b.close()


@pytest.mark.parametrize("budgets", [(0, 1, 5000), (137, 10**5, 2), (3000, 3001, 2999)])
def test_rows_budget_back_pressure(ctx, budgets):
    """hbt_set_rows_budget: the SAMPLE whose row the sink cannot take ends that CPU's consume() with
    -ENOSPC in front of it (the reference with a full output ring); the rest of the window, decoded
    after the sink was drained, starts by re-seeding on that SAMPLE.  Cuts land in first tiles,
    at segment seams and deep inside windows."""
    ds, E = 1 << 20, 6
    desc = hbt.sampling_desc(E)
    rings = []
    for cpu in range(3):
        s = tg.group_sample_stream(700 + cpu, 9000, E, lost_every=(0, 97, 1500)[cpu])
        tail = tg.random_tail(50 + cpu, ds)
        ring, head = ring_np(s, ds, tail)
        rings.append((ring, tail, head))
    ctx.ring_configure(3, ds)
    ctx.reset_state(123)
    ctx.set_rows_budget(budgets)
    try:
        res = gpu_decode(ctx, rings, desc, ds, reset=False)
        wants = []
        for cpu, (ring, tail, head) in enumerate(rings):
            w = orc.consume(ring, tail, head, desc, 123, np.zeros(E, np.uint64), rows_budget=budgets[cpu])
            check_cpu(ctx, cpu, res[cpu], w, desc)
            assert (w["ret"] == -28) == (budgets[cpu] < 8000)
            wants.append(w)
        check_rows_all(ctx, wants, E)
        # the sinks were drained: the rest of every window
        ctx.set_rows_budget(None)
        rest = [(ring, wants[c]["new_tail"], head) for c, (ring, tail, head) in enumerate(rings)]
        res2 = gpu_decode(ctx, rest, desc, ds, reset=False)
        for cpu, (ring, tail, head) in enumerate(rest):
            w0 = wants[cpu]
            w = orc.consume(ring, tail, head, desc, w0["carry_tstamp"], w0["carry_last"], w0["counters"],
                            w0["num_unknown"])
            check_cpu(ctx, cpu, res2[cpu], w, desc)
    finally:
        ctx.set_rows_budget(None)


def test_mixed_stream_deltas(ctx):
    ds, E = 1 << 21, 8
    mix = (("sample", 0.6), ("switch", 0.2), ("comm", 0.05), ("mmap", 0.05), ("lost", 0.02),
           ("fork", 0.04), ("exit", 0.04))
    rings = []
    for cpu in range(3):
        s = tg.mixed_stream(43 + cpu, 20000, E, mix=mix)
        tail = tg.random_tail(10 + cpu, ds)
        ring, head = ring_np(s, ds, tail)
        rings.append((ring, tail, head))
    desc = hbt.sampling_desc(E, handled=hbt.MODE_MIXED)
    res = gpu_decode(ctx, rings, desc, ds)
    for cpu, (ring, tail, head) in enumerate(rings):
        check_cpu(ctx, cpu, res[cpu], orc.consume(ring, tail, head, desc), desc)


# ---------------------------------------------------------------------------
# (cpu,pid,ip) histogram (K6)
# ---------------------------------------------------------------------------
@pytest.mark.parametrize("words,ds,n_cpus,n_pids,n_ips", [
    (5, 1 << 18, 3, 64, 256), (6, 1 << 20, 2, 512, 4096), (16, 1 << 20, 2, 16, 16),
    (5, 1 << 22, 6, 4096, 65536)])
def test_ip_histogram(ctx, words, ds, n_cpus, n_pids, n_ips):
    desc = hbt.ip_desc(8 * words)
    rings, want_hist, wants = [], {}, []
    for cpu in range(n_cpus):
        n = (ds - 8192 - 64 * cpu) // (8 * words)
        s = tg.ip_sample_stream(51 + cpu, n, cpu=cpu, sample_words=words, n_pids=n_pids,
                                n_ips=n_ips, lost_every=997)
        tail = tg.random_tail(12 + cpu, ds)
        ring, head = ring_np(s, ds, tail)
        rings.append((ring, tail, head))
        w = orc.consume(ring, tail, head, desc, cpu=cpu)
        want_hist.update(w["hist"])
        wants.append(w)
    res = gpu_decode(ctx, rings, desc, ds, hist_slots=1 << 22)
    for cpu in range(n_cpus):
        check_cpu(ctx, cpu, res[cpu], wants[cpu], desc)
    got = orc.hist_to_dict(ctx.hist_fetch())
    assert len(got) == len(want_hist)
    assert got == want_hist
    # the same table as 16-byte entries behind a per-CPU offset table (hbt_hist_fetch_packed), whole and
    # for a sub-range of the CPUs
    ent, off = ctx.hist_fetch_packed()
    assert int(off[0]) == 0 and int(off[-1]) == len(ent) == len(want_hist)
    packed = {(cpu, int(e["pid"]), int(e["ip"])): int(e["count"])
              for cpu in range(n_cpus) for e in ent[int(off[cpu]):int(off[cpu + 1])]}
    assert packed == want_hist
    if n_cpus > 2:
        ent, off = ctx.hist_fetch_packed(1, n_cpus - 2)
        sub = {(1 + i, int(e["pid"]), int(e["ip"])): int(e["count"])
               for i in range(n_cpus - 2) for e in ent[int(off[i]):int(off[i + 1])]}
        assert sub == {k: v for k, v in want_hist.items() if 1 <= k[0] < n_cpus - 1}
    with pytest.raises(hbt.HbtError):
        ctx.hist_fetch_packed(0, n_cpus, out=np.zeros(3, dtype=hbt.HIST_PACKED_DTYPE))   # -ENOSPC
    # a second identical batch accumulates
    for cpu, (ring, tail, head) in enumerate(rings):
        ctx.submit_ring_image(cpu, ring, tail, head)
    ctx.decode(desc)
    ctx.fetch_results()
    got2 = orc.hist_to_dict(ctx.hist_fetch())
    assert got2 == {k: 2 * v for k, v in want_hist.items()}


def test_ip_histogram_zero_key_and_error(ctx):
    """Key (ip=0,pid=0) (kept outside the table) and a bad-size SAMPLE in the
    middle: everything after it must be taken back out of the shared table."""
    ds, words = 1 << 20, 5
    n = 20000
    s = tg.ip_sample_stream(7, n, cpu=0, n_pids=8, n_ips=8)
    s = s.clone()
    s[1::5][:50] = 0       # ip = 0
    s[2::5][:50] = 0       # pid = tid = 0
    bad_at = 12345
    s[5 * bad_at] = tg.header(tg.REC_SAMPLE, 48)  # size != ip_sample_bytes
    tail = tg.random_tail(3, ds)
    ring, head = ring_np(s, ds, tail)
    desc = hbt.ip_desc(40)
    want = orc.consume(ring, tail, head, desc)
    assert want["ret"] == -1 and want["n_records"] == bad_at
    res = gpu_decode(ctx, [(ring, tail, head)], desc, ds, hist_slots=1 << 16)
    check_cpu(ctx, 0, res[0], want, desc)
    assert orc.hist_to_dict(ctx.hist_fetch()) == want["hist"]
    assert (0, 0, 0) in want["hist"]


# ---------------------------------------------------------------------------
# hostile / degenerate inputs: exactness must not depend on the guesses
# ---------------------------------------------------------------------------
def test_payload_that_looks_like_headers(ctx):
    """Counter values crafted to be plausible perf_event_headers whose chains even
    validate: the guesses go wrong, the verification catches them, results stay exact."""
    ds, E = 1 << 19, 8
    s = tg.group_sample_stream(5, 4000, E).clone()
    W = 5 + E
    fake = tg.header(tg.REC_SAMPLE, 8 * W)  # a value equal to a real SAMPLE header
    s[7::W] = fake                           # value[2] of every sample
    s[9::W] = tg.header(tg.REC_LOST, 24)
    # a LOST record in front shifts every SAMPLE off the positions an equal-sized-records stream
    # would have them at, so the segments cannot lean on the host's entry expectation: they scan
    import torch
    s = torch.cat([torch.tensor([tg.header(tg.REC_LOST, 24), 7, 3], dtype=torch.int64), s])
    tail = tg.random_tail(77, ds)
    ring, head = ring_np(s, ds, tail, fill=0)
    desc = hbt.sampling_desc(E)
    clean = tg.group_sample_stream(6, 4000, E)
    ring_c, head_c = ring_np(clean, ds, tail)
    ctx.ring_configure(3, ds)
    ctx.set_geometry(16, 1)          # many short segments: many entry guesses to get wrong
    try:
        res = gpu_decode(ctx, [(ring, tail, head), (ring_c, tail, head_c), (ring, tail, head)], desc, ds, reset=False)
    finally:
        ctx.set_geometry(0, 0)
    want = orc.consume(ring, tail, head, desc)
    check_cpu(ctx, 0, res[0], want, desc)
    check_cpu(ctx, 1, res[1], orc.consume(ring_c, tail, head_c, desc), desc)
    check_cpu(ctx, 2, res[2], want, desc)
    # wrong guesses cost re-runs, but only of the segments that were wrong: never the clean CPU's
    # (an attempt re-decodes at most one segment per hostile CPU)
    if ctx.kernel == hbt.KERNEL_WALK:
        # k_walk repairs wrong entry guesses inside the launch (look-back at the segment in front; a walk
        # that stopped at an error publishes its exit only after its own entry is settled, so even this
        # stream, whose fake headers form a chain of their own, is repaired down the chain)
        assert ctx.slow_tiles > 0 and ctx.reruns == 0, (ctx.slow_tiles, ctx.reruns)
    else:
        assert ctx.reruns > 0
        assert ctx.rerun_segments <= 2 * ctx.reruns


def test_walk_repairs_a_wrong_entry_guess_in_place(built):
    """One fake header right behind a segment boundary, inside the SAMPLE that straddles it, whose size
    leads to the next real record: the segment's guess takes it for its first record.  k_walk's look-back
    (exit of the segment in front) finds the mistake and walks the segment again inside the same launch:
    exact results, no re-run from the host."""
    ds, E = 1 << 19, 8
    S = 40 + 8 * E
    ctx = hbt.Context(0)
    try:
        ctx.set_decode_kernel(hbt.KERNEL_WALK)
        s = tg.group_sample_stream(11, 4000, E).clone()
        tail = tg.random_tail(78, ds)
        head = tail + 8 * len(s)
        # segment geometry of hbt_set_geometry(16, 1): per tiles of 16 KiB from the tail's tile
        t0, t1 = tail // 16384, (head + 16383) // 16384
        per = max(-(-(t1 - t0) // 16), 1)
        planted = 0
        for k in range(1, 16):
            B = (t0 + k * per) * 16384
            if B >= head - 4 * S:
                break
            i = (B - tail) // S                 # the SAMPLE that straddles the boundary
            p0 = tail + i * S
            q = max(B, p0 + 40 + 8)             # a value word at or behind the boundary
            q = (q + 7) & ~7
            if q >= p0 + S - 8 or q < p0 + 40:
                continue
            s[(q - tail) // 8] = tg.header(tg.REC_SAMPLE, p0 + S - q)
            planted += 1
        assert planted >= 4
        ring, head2 = ring_np(s, ds, tail, fill=0)
        assert head2 == head
        desc = hbt.sampling_desc(E)
        ctx.ring_configure(1, ds)
        ctx.set_geometry(16, 1)
        res = gpu_decode(ctx, [(ring, tail, head)], desc, ds, reset=False)
        check_cpu(ctx, 0, res[0], orc.consume(ring, tail, head, desc), desc)
        assert ctx.slow_tiles >= planted - 1 and ctx.reruns == 0
    finally:
        ctx.close()


def test_unknown_types_and_giant_records(ctx):
    """Types the guess filter rejects (>= 64) and records larger than a tile take
    the sequential path; consume() just skips them by size."""
    ds = 1 << 20
    words = []
    rng = np.random.default_rng(1)
    total = 0
    while total < (ds - 70000) // 8:
        k = rng.integers(0, 10)
        if k == 0:
            w = int(rng.integers(2100, 8191))          # 16.8..65.5 KB record, type 1000
            rec = [1000 | (8 * w) << 48] + [int(x) for x in rng.integers(0, 2**40, w - 1)]
        elif k < 4:
            w = int(rng.integers(1, 40))
            rec = [int(rng.integers(64, 5000)) | (8 * w) << 48] + [0] * (w - 1)
        else:
            rec = [9 | (8 * 7) << 48, 1000 + total, 2, 5, 5, total, 2 * total]
        words += rec
        total += len(rec)
    import torch
    s = torch.tensor(np.array(words, dtype=np.uint64).view(np.int64))
    tail = tg.random_tail(8, ds)
    ring, head = ring_np(s, ds, tail)
    desc = hbt.sampling_desc(2)
    res = gpu_decode(ctx, [(ring, tail, head)], desc, ds)
    check_cpu(ctx, 0, res[0], orc.consume(ring, tail, head, desc), desc)


@pytest.mark.parametrize("mode", [hbt.MODE_SAMPLING, hbt.MODE_MIXED])
def test_tiny_records(ctx, mode):
    """Runs of 8- and 16-byte records (THROTTLE / UNTHROTTLE are counted in every mode, type 70 is only
    skipped): up to 1024 record starts per 8 KiB, the most k_walk's per-round record list and the 8-bit
    per-lane type counters have to take, with ordinary samples in between."""
    ds = 1 << 19
    rng = np.random.default_rng(3)
    words, total = [], 0
    while total < (ds - 9000) // 8:
        k = int(rng.integers(0, 6))
        if k == 0:      # a long run of 8-byte THROTTLE records
            rec = [tg.header(tg.REC_THROTTLE, 8)] * int(rng.integers(200, 3000))
        elif k == 1:    # 16-byte UNTHROTTLE records
            rec = [tg.header(tg.REC_UNTHROTTLE, 16), 5] * int(rng.integers(50, 800))
        elif k == 2:    # 8-byte records of a type nobody handles
            rec = [70 | (8 << 48)] * int(rng.integers(10, 600))
        else:
            rec = [tg.header(tg.REC_SAMPLE, 56), 1000 + total, 2, 5, 5, total, 2 * total]
        words += rec
        total += len(rec)
    import torch
    s = torch.tensor(np.array(words, dtype=np.uint64).view(np.int64))
    tail = tg.random_tail(21, ds)
    ring, head = ring_np(s, ds, tail)
    for desc in (hbt.sampling_desc(2), hbt.count_desc(mode)):
        res = gpu_decode(ctx, [(ring, tail, head)], desc, ds)
        check_cpu(ctx, 0, res[0], orc.consume(ring, tail, head, desc), desc)


def test_zero_size_record_stops(ctx):
    ds, E = 1 << 17, 8
    s = tg.group_sample_stream(5, 800, E).clone()
    s[(5 + E) * 600] = 9  # SAMPLE header with size 0
    ring, head = ring_np(s, ds, 0)
    desc = hbt.sampling_desc(E)
    res = gpu_decode(ctx, [(ring, 0, head)], desc, ds)
    # (the reference itself would never return from this stream: documented deviation,
    # so the restatement alone is the checker here)
    want = orc.port_consume(ring, 0, head, desc)
    assert want["ret"] == -74 and want["n_records"] == 600
    check_cpu(ctx, 0, res[0], want, desc)


@pytest.mark.parametrize("ds", [4096, 8192, 16384, 32768])
def test_small_rings_and_empty_windows(ctx, ds):
    E = 2
    desc = hbt.sampling_desc(E)
    rings = []
    for cpu in range(4):
        n = 0 if cpu == 1 else (ds - 256) // (8 * (5 + E) + 3)
        s = tg.group_sample_stream(cpu + 1, max(n, 1), E, lost_every=9)
        if n == 0:
            s = s[:0]
        tail = 5 * ds + 8 * 13 * cpu
        ring, head = ring_np(s, ds, tail)
        rings.append((ring, tail, head))
    res = gpu_decode(ctx, rings, desc, ds)
    wants = [orc.consume(ring, tail, head, desc) for ring, tail, head in rings]
    for cpu in range(len(rings)):
        check_cpu(ctx, cpu, res[cpu], wants[cpu], desc)
    assert int(res[1]["ret"]) == -61 and int(res[1]["n_records"]) == 0
    check_rows_all(ctx, wants, E)   # an empty CPU in the middle


def test_window_overrunning_head(ctx):
    """head in the middle of a record: consume() still takes the record that
    starts before head and leaves tail past head (CpuEventsGroup.h:1348,1504)."""
    ds, E = 1 << 18, 8
    s = tg.group_sample_stream(3, 2000, E)
    ring, head = ring_np(s, ds, 64)
    desc = hbt.sampling_desc(E)
    for cut in (8, 16, 96):
        res = gpu_decode_full_images(ctx, [(ring, 64, head - cut)], desc, ds)
        want = orc.consume(ring, 64, head - cut, desc)
        assert want["new_tail"] > head - cut or cut % 104 == 0
        check_cpu(ctx, 0, res[0], want, desc)


def test_device_resident_rings(ctx):
    """hbt_submit_ring_dev: caller-owned device images (the benchmark's path)."""
    import torch
    ds, E, n_cpus = 1 << 20, 8, 8
    desc = hbt.sampling_desc(E)
    imgs = torch.zeros((n_cpus, ds // 8), dtype=torch.int64, device="cuda")
    meta = []
    for cpu in range(n_cpus):
        s = tg.group_sample_stream(300 + cpu, 9000, E, device="cuda", lost_every=250)
        tail = tg.random_tail(cpu, ds)
        ring, head = tg.place_in_ring(s, ds, tail, fill_seed=5)
        imgs[cpu] = ring
        meta.append((tail, head))
    torch.cuda.synchronize()
    ctx.ring_configure(n_cpus, 0)
    ctx.reset_state(hbt.INVALID_TS)
    for cpu, (tail, head) in enumerate(meta):
        ctx.submit_ring_dev(cpu, imgs[cpu], ds, tail, head)
    ctx.decode(desc)
    res = ctx.fetch_results()
    host = imgs.cpu().numpy().view(np.uint64)
    for cpu, (tail, head) in enumerate(meta):
        check_cpu(ctx, cpu, res[cpu], orc.consume(host[cpu], tail, head, desc), desc)


# ---------------------------------------------------------------------------
# API behaviour around the kernels
# ---------------------------------------------------------------------------
def test_hist_capacity_overflow_is_loud(ctx):
    """Lists too small for one decode: -ENOSPC from fetch_results, never a silent drop."""
    import errno
    ds = 1 << 20
    n = (ds - 8192) // 40
    s = tg.ip_sample_stream(9, n, cpu=0, n_pids=4096, n_ips=65536)
    ring, head = ring_np(s, ds, 0)
    ctx.ring_configure(1, ds)
    ctx.reset_state(hbt.INVALID_TS)
    ctx.hist_configure(4096)  # far fewer entries than samples
    ctx.submit_ring_image(0, ring, 0, head)
    ctx.decode(hbt.ip_desc(40))
    with pytest.raises(hbt.HbtError) as e:
        ctx.fetch_results()
    assert e.value.code == -errno.ENOSPC
    # a sufficient capacity works, and clear() empties the lists
    ctx.hist_configure(1 << 18)
    ctx.submit_ring_image(0, ring, 0, head)
    ctx.decode(hbt.ip_desc(40))
    ctx.fetch_results()
    want = orc.consume(ring, 0, head, hbt.ip_desc(40))["hist"]
    assert orc.hist_to_dict(ctx.hist_fetch()) == want
    ctx.hist_clear()
    assert ctx.hist_size() == 0


def test_hist_must_be_reconfigured_after_cpu_count_change(built):
    """ring_configure(8) -> hist_configure -> ring_configure(64): the lists were laid out for 8
    CPUs; a histogram decode is refused (-EINVAL) until hbt_hist_configure is called again."""
    import errno
    ds = 1 << 16
    ctx = hbt.Context(0)
    try:
        ctx.ring_configure(8, ds)
        ctx.hist_configure(8 << 13)
        ctx.ring_configure(64, ds)
        rings = []
        for cpu in range(64):
            s = tg.ip_sample_stream(600 + cpu, 1200, cpu=cpu, n_pids=8, n_ips=32)
            ring, head = ring_np(s, ds, 0)
            rings.append((ring, 0, head))
            ctx.submit_ring_image(cpu, ring, 0, head)
        with pytest.raises(hbt.HbtError) as e:
            ctx.decode(hbt.ip_desc(40))
        assert e.value.code == -errno.EINVAL
        ctx.hist_configure(64 << 13)
        for cpu, (ring, tail, head) in enumerate(rings):
            ctx.submit_ring_image(cpu, ring, tail, head)
        ctx.decode(hbt.ip_desc(40))
        ctx.fetch_results()
        want = {}
        for cpu, (ring, tail, head) in enumerate(rings):
            want.update(orc.consume(ring, tail, head, hbt.ip_desc(40), cpu=cpu)["hist"])
        assert orc.hist_to_dict(ctx.hist_fetch()) == want
    finally:
        ctx.close()


def test_heavy_hitter_key(ctx):
    """One key carries 90 % of the samples: the segment cache must absorb it (a list
    holds 2048 entries) and the count must still be exact."""
    ds = 1 << 22
    n = (ds - 8192) // 40
    s = tg.ip_sample_stream(10, n, cpu=0, n_pids=64, n_ips=4096).clone()
    hot = tg.rand_uniform(77, n) < 0.9
    s.view(n, 5)[hot, 1] = 0x401000
    s.view(n, 5)[hot, 2] = 4242 | (4242 << 32)
    tail = tg.random_tail(4, ds)
    ring, head = ring_np(s, ds, tail)
    desc = hbt.ip_desc(40)
    res = gpu_decode(ctx, [(ring, tail, head)], desc, ds, hist_slots=1 << 18)
    want = orc.consume(ring, tail, head, desc)
    check_cpu(ctx, 0, res[0], want, desc)
    got = orc.hist_to_dict(ctx.hist_fetch())
    assert got == want["hist"]
    assert got[(0, 4242, 0x401000)] > 0.85 * n


def test_batch_submit_and_device_views(ctx):
    import ctypes as C
    import torch
    ds, E, n_cpus = 1 << 20, 4, 3
    desc = hbt.sampling_desc(E)
    imgs = torch.zeros((n_cpus, ds // 8), dtype=torch.int64, device="cuda")
    tails, heads = [], []
    for cpu in range(n_cpus):
        s = tg.group_sample_stream(500 + cpu, 9000, E, device="cuda", lost_every=777)
        tail = tg.random_tail(cpu + 20, ds)
        ring, head = tg.place_in_ring(s, ds, tail, fill_seed=5)
        imgs[cpu] = ring
        tails.append(tail)
        heads.append(head)
    torch.cuda.synchronize()
    ctx.ring_configure(n_cpus, 0)
    ctx.reset_state(hbt.INVALID_TS)
    bases = np.array([imgs[c].data_ptr() for c in range(n_cpus)], dtype=np.uint64)
    ctx.submit_rings_dev(bases, ds, np.array(tails, np.uint64), np.array(heads, np.uint64))
    ctx.decode(desc)
    res = ctx.fetch_results()
    host = imgs.cpu().numpy().view(np.uint64)
    for cpu in range(n_cpus):
        want = orc.consume(host[cpu], tails[cpu], heads[cpu], desc)
        check_cpu(ctx, cpu, res[cpu], want, desc)
        # device view of the same rows: slabs in stream order
        k = ctx.lib.hbt_rows_max_slabs(ctx.h)
        ptrs = (C.c_void_p * k)()
        ns = (C.c_uint64 * k)()
        n_slabs = C.c_uint32()
        assert ctx.lib.hbt_rows_slabs_dev(ctx.h, cpu, ptrs, ns, C.byref(n_slabs)) == 0
        assert sum(ns[i] for i in range(n_slabs.value)) == want["n_rows"]


def test_two_contexts_interleaved(built):
    """Two contexts (own streams) with decodes in flight at the same time, the way bench.py's
    end-to-end leg pipelines CPU groups over PCIe: results are those of separate decodes."""
    ds, E = 1 << 19, 4
    a, b = hbt.Context(0), hbt.Context(0)
    try:
        jobs = []
        for ctx, seed, desc, base in ((a, 71, hbt.sampling_desc(E), 0), (b, 72, hbt.ip_desc(40), 8)):
            rings = []
            for c in range(3):
                if desc.aggregate == hbt.AGG_COUNT_DELTAS:
                    s = tg.group_sample_stream(seed * 10 + c, 3000, E, lost_every=301)
                else:
                    s = tg.ip_sample_stream(seed * 10 + c, 9000, cpu=base + c, n_pids=32, n_ips=128)
                tail = tg.random_tail(seed + c, ds)
                ring, head = ring_np(s, ds, tail)
                rings.append((ring, tail, head))
            ctx.ring_configure(3, ds)
            ctx.set_cpu_base(base)
            ctx.reset_state(hbt.INVALID_TS)
            if desc.aggregate == hbt.AGG_IP_HISTOGRAM:
                ctx.hist_configure(3 << 16)
            jobs.append((ctx, rings, desc, base))
        # everything queued before anything is fetched
        for ctx, rings, desc, _ in jobs:
            for cpu, (ring, tail, head) in enumerate(rings):
                ctx.submit_ring_image(cpu, ring, tail, head)
            ctx.decode(desc)
        for ctx, rings, desc, base in reversed(jobs):
            res = ctx.fetch_results()
            hist = orc.hist_to_dict(ctx.hist_fetch()) if desc.aggregate == hbt.AGG_IP_HISTOGRAM else None
            want_hist = {}
            for cpu, (ring, tail, head) in enumerate(rings):
                want = orc.consume(ring, tail, head, desc, cpu=base + cpu)
                check_cpu(ctx, cpu, res[cpu], want, desc)
                if want.get("hist"):
                    want_hist.update(want["hist"])
            if hist is not None:
                assert hist == want_hist
    finally:
        a.close()
        b.close()


def test_max_cpus_decode(built):
    """HBT_MAX_CPUS rings in one decode (the reference's kMaxCpus, System.h:195): every CPU's
    counters, rows and carry against the oracle."""
    ds, E, n_cpus = 1 << 16, 3, hbt.MAX_CPUS
    ctx = hbt.Context(0)
    try:
        rings = []
        for c in range(n_cpus):
            s = tg.group_sample_stream(9000 + c, 200 + (c % 7) * 40, E, lost_every=53 if c % 3 else 0)
            tail = tg.random_tail(c, ds)
            ring, head = ring_np(s, ds, tail)
            rings.append((ring, tail, head))
        desc = hbt.sampling_desc(E)
        res = gpu_decode(ctx, rings, desc, ds)
        for c in (0, 1, 2, 255, 256, 300, 510, 511):
            check_cpu(ctx, c, res[c], orc.consume(*rings[c], desc), desc)
        want_rec = [orc.consume(*rings[c], desc)["n_records"] for c in range(n_cpus)]
        assert [int(r) for r in res["n_records"]] == want_rec
    finally:
        ctx.close()
