"""Context-switch records -> tagstack::Event stream (HBT_AGG_SWITCH_EVENTS, SURVEY 8d config 4
"switch-event table" / 8f rank 2).

CPU part: the C restatement (oracle/hbt_oracle.c) + the test-side vid mapper are pinned to the
REAL CpuThreadSwitchGenerator (oracle/_ref, PerCpuThreadSwitchGenerator.h:168-454) on streams
its handlers accept.  GPU part: the CUDA path through the C-ABI against the C restatement,
bit-exact, on the same seeded inputs."""
import numpy as np
import pytest

import dynolog_1_b200 as hbt
from dynolog_1_b200 import tracegen as tg
import _oracle as orc


def ring_np(stream, ds, tail, fill=77):
    ring, head = tg.place_in_ring(stream, ds, tail, fill_seed=fill)
    return ring.numpy().view(np.uint64), head


def consistent_rings(seed, n_cpus, n_records, ds, lost_every):
    rings, tails, heads = [], [], []
    for c in range(n_cpus):
        s = tg.ctx_switch_stream_consistent(seed * 100 + c, n_records,
                                            lost_every=lost_every if c != 1 else 0,
                                            first_tid=1000 + 10000 * c)
        tail = 5 * ds + ds - 2048 - 8 * c  # the window wraps the ring end
        ring, head = ring_np(s, ds, tail)
        rings.append(ring)
        tails.append(tail)
        heads.append(head)
    return np.stack(rings), tails, heads


def same_events(a, b):
    return len(a) == len(b) and bool((a == b).all())


# ---------------------------------------------------------------------------
# oracle vs the reference's own generator
# ---------------------------------------------------------------------------
@pytest.mark.parametrize("seed,lost_every", [(1, 57), (2, 11), (3, 0), (4, 2)])
def test_oracle_vs_reference_generator(built, seed, lost_every):
    if not orc.have_ref():
        pytest.skip("oracle/_ref not built (reference tree absent)")
    ds = 1 << 16
    rings, tails, heads = consistent_rings(seed, 3, 700, ds, lost_every)
    d = hbt.switch_desc(tid_level=2)
    ref = orc.ref_switch_events(rings, tails, heads, tid_level=2, cpu_base=4)
    port = [orc.port_switch_events(rings[c], tails[c], heads[c], d, cpu=4 + c) for c in range(3)]
    mapped = orc.map_tids_to_vids([p["events"] for p in port])
    for c in range(3):
        for k in ("ret", "n_records", "new_tail", "counters", "carry_tstamp", "num_unknown"):
            assert ref[c][k] == port[c][k], (c, k)
        assert same_events(ref[c]["events"], mapped[c]), c
        # what the pseudo-events are for: one per COMM record, nothing else differs in count
        assert len(port[c]["events"]) - len(mapped[c]) == port[c]["counters"][hbt.CNT_COMM]


@pytest.mark.parametrize("seed,mix", [(11, tg.CTX_MIX), (12, (("switch", 0.5), ("lost", 0.2), ("exit", 0.15), ("fork", 0.15))),
                                      (13, (("comm", 0.5), ("exit", 0.3), ("fork", 0.2)))])
def test_oracle_vs_reference_generator_random_tids(built, seed, mix):
    """The vectorised stream (random tids from a small pool, tid 0 included, exits of threads
    never seen before, COMM of unknown threads): the reference's generator accepts it, and its
    vid-tagged events equal the restatement's mapped ones."""
    if not orc.have_ref():
        pytest.skip("oracle/_ref not built (reference tree absent)")
    ds = 1 << 17
    rings, tails, heads = [], [], []
    for c in range(2):
        s = tg.ctx_switch_stream(seed * 10 + c, 2500, mix=mix, n_tids=48, fork_tid_base=100000 * c)
        tail = tg.random_tail(seed + c, ds)
        ring, head = ring_np(s, ds, tail)
        rings.append(ring)
        tails.append(tail)
        heads.append(head)
    rings = np.stack(rings)
    d = hbt.switch_desc(tid_level=1)
    ref = orc.ref_switch_events(rings, tails, heads, tid_level=1)
    port = [orc.port_switch_events(rings[c], tails[c], heads[c], d, cpu=c) for c in range(2)]
    mapped = orc.map_tids_to_vids([p["events"] for p in port])
    for c in range(2):
        for k in ("ret", "n_records", "new_tail", "counters", "carry_tstamp", "num_unknown"):
            assert ref[c][k] == port[c][k], (c, k)
        assert same_events(ref[c]["events"], mapped[c]), c


def test_oracle_vs_reference_carry_and_max_records(built):
    if not orc.have_ref():
        pytest.skip("oracle/_ref not built (reference tree absent)")
    ds = 1 << 15
    rings, tails, heads = consistent_rings(9, 1, 400, ds, 13)
    # generator state carried in: inside a write-error period, last_tstamp_ set
    for max_records in (hbt.DRAIN, 123, 1, 0):
        d = hbt.switch_desc(tid_level=1, max_records=max_records)
        ref = orc.ref_switch_events(rings, tails, heads, tid_level=1, max_records=max_records,
                                    last_tstamp=[777], in_write_errors=[1])[0]
        port = orc.port_switch_events(rings[0], tails[0], heads[0], d, last_tstamp=777, in_write_errors=1)
        for k in ("ret", "n_records", "new_tail", "counters", "carry_tstamp", "num_unknown"):
            assert ref[k] == port[k], (max_records, k)
        assert same_events(ref["events"], orc.map_tids_to_vids([port["events"]])[0]), max_records
        if max_records:
            assert port["events"][0]["type"] == hbt.SWEV_WRITE_ERRORS_END


def test_oracle_event_semantics(built):
    """Hand-written stream: the exact events of PerCpuThreadSwitchGenerator.h:228-412."""
    def hdr(t, words, misc=0):
        return t | (misc << 32) | ((8 * words) << 48)
    OUT, PRE = tg.MISC_SWITCH_OUT, tg.MISC_SWITCH_OUT | tg.MISC_SWITCH_OUT_PREEMPT
    w = []
    w += [hdr(15, 4), 0, 40 | (41 << 32), 100]                 # switch in tid 41 @100
    w += [hdr(15, 4, OUT), 0, 40 | (41 << 32), 110]            # switch out (yield) @110
    w += [hdr(2, 5), 7, 3, 0, 150]                             # LOST(3) @150
    w += [hdr(2, 5), 7, 2, 0, 160]                             # LOST(2) @160, back to back
    w += [hdr(15, 4, PRE), 0, 40 | (41 << 32), 170]            # switch out (preempt) @170
    w += [hdr(7, 6), 50 | (1 << 32), 51 | (1 << 32), 180, 0, 181]   # FORK tid 51 @180
    w += [hdr(3, 5), 50 | (51 << 32), 0x6161, 50 | (51 << 32), 190]  # COMM tid 51 @190
    w += [hdr(4, 6), 50 | (1 << 32), 51 | (1 << 32), 200, 0, 201]   # EXIT tid 51 @200
    s = np.array(w, dtype=np.uint64)
    ds = 4096
    ring = np.zeros(ds // 8, dtype=np.uint64)
    ring[:len(s)] = s
    got = orc.port_switch_events(ring, 0, 8 * len(s), hbt.switch_desc(tid_level=3), cpu=6)
    ev = [(int(e["tstamp"]), int(e["type"]), int(e["level"]), int(e["comp_unit_id"]), int(e["tag"]))
          for e in got["events"]]
    NA, INV = hbt.TAG_NA, hbt.LEVEL_INVALID
    assert ev == [
        (100, hbt.SWEV_SWITCH_IN, 3, 6, 41),
        (110, hbt.SWEV_SWITCH_OUT_YIELD, 3, 6, NA),
        (111, hbt.SWEV_WRITE_ERRORS_START, INV, 6, NA),        # last_tstamp_ + 1
        (160, hbt.SWEV_WRITE_ERRORS_END, INV, 6, NA),          # pendingRecordLost_ of the 2nd LOST
        (161, hbt.SWEV_WRITE_ERRORS_START, INV, 6, NA),
        (170, hbt.SWEV_WRITE_ERRORS_END, INV, 6, NA),
        (170, hbt.SWEV_SWITCH_OUT_PREEMPT, 3, 6, NA),
        (180, hbt.SWEV_THREAD_CREATION, 3, 6, 51),
        (190, hbt.SWEV_COMM, 3, 6, 51),
        (200, hbt.SWEV_THREAD_DESTRUCTION, 3, 6, 51),
    ]
    assert got["ret"] == -61 and got["counters"][hbt.CNT_LOST] == 5 and got["carry_tstamp"] == 200
    # a SWITCH record shorter than its struct: refused with -EPERM in front of the record
    bad = s.copy()
    bad[4] = hdr(15, 3, OUT)
    got = orc.port_switch_events(np.concatenate([bad, np.zeros(64, np.uint64)])[:512], 0, 8 * len(s),
                                 hbt.switch_desc())
    assert got["ret"] == -1 and got["n_records"] == 1 and len(got["events"]) == 1


def test_combinator_restatement_known_answers(built):
    """hbt/src/tagstack/tests/StreamTest.cpp:61-130 (BinaryCombinator): e1..e6 come out in
    that order; equal timestamps are ordered by comp_unit_id."""
    def ev(ts, t, lvl, tag, cu):
        a = np.zeros(1, dtype=hbt.SWITCH_EVENT_DTYPE)
        a[0] = (ts, t, lvl, cu, tag)
        return a
    NA = hbt.TAG_NA
    e1 = ev(20, hbt.SWEV_SWITCH_IN, 0, 2, 2)
    e2 = ev(21, hbt.SWEV_SWITCH_OUT_PREEMPT, 0, NA, 2)
    e3 = ev(24, hbt.SWEV_SWITCH_IN, 1, 2, 1)
    e4 = ev(30, hbt.SWEV_SWITCH_OUT_PREEMPT, 1, NA, 1)
    e5 = ev(30, hbt.SWEV_SWITCH_IN, 1, 5, 2)
    e6 = ev(31, hbt.SWEV_SWITCH_OUT_PREEMPT, 1, NA, 1)
    got = orc.combinator_merge([np.concatenate([e3, e4, e6]), np.concatenate([e1, e2, e5])])
    assert same_events(got, np.concatenate([e1, e2, e3, e4, e5, e6]))
    # five streams (odd node at two levels) against a stable sort by (tstamp, comp_unit_id)
    rng = np.random.default_rng(5)
    streams = []
    for cu in range(5):
        n = int(rng.integers(0, 40))
        a = np.zeros(n, dtype=hbt.SWITCH_EVENT_DTYPE)
        a["tstamp"] = np.sort(rng.integers(0, 50, n))
        a["comp_unit_id"] = cu
        a["tag"] = np.arange(n)
        streams.append(a)
    got = orc.combinator_merge(streams)
    allev = np.concatenate(streams)
    want = allev[np.lexsort((allev["comp_unit_id"], allev["tstamp"]))]
    assert same_events(got, want)


# ---------------------------------------------------------------------------
# CUDA path vs oracle
# ---------------------------------------------------------------------------
@pytest.fixture(scope="module", params=[hbt.KERNEL_TILES, hbt.KERNEL_WALK], ids=["tiles", "walk"])
def ctx(built, request):
    """Every test runs under both decode kernels (the choice only matters where k_walk covers the
    aggregation; results must never depend on it)."""
    c = hbt.Context(0)
    c.set_decode_kernel(request.param)
    yield c
    c.close()


def gpu_switch(ctx, rings, desc, ds, reset=True, cpu_base=0):
    if reset:
        ctx.ring_configure(len(rings), ds)
        ctx.reset_state(hbt.INVALID_TS)
    ctx.set_cpu_base(cpu_base)
    for cpu, (ring, tail, head) in enumerate(rings):
        ctx.submit_ring_image(cpu, ring, tail, head)
    ctx.decode(desc)
    res = ctx.fetch_results()
    ev = [ctx.fetch_switch_events(c, int(res[c]["n_rows"])) for c in range(len(rings))]
    return res, ev


def check(res, ev, want, c):
    assert int(res[c]["ret"]) == want["ret"], (c, int(res[c]["ret"]), want["ret"])
    assert int(res[c]["n_records"]) == want["n_records"], c
    assert int(res[c]["new_tail"]) == want["new_tail"], c
    assert [int(x) for x in res[c]["counters"]] == want["counters"], c
    assert int(res[c]["carry_tstamp"]) == want["carry_tstamp"], c
    assert int(res[c]["num_unknown_count_interval"]) == want["num_unknown"], c
    assert int(res[c]["n_rows"]) == len(want["events"]), c
    assert same_events(ev[c], want["events"]), f"events differ on cpu {c}"


@pytest.mark.gpu
@pytest.mark.parametrize("n_records,ds,n_cpus,mix", [
    (3000, 1 << 17, 5, tg.CTX_MIX),
    (60000, 1 << 21, 3, tg.CTX_MIX),
    (20000, 1 << 20, 2, (("switch", 0.5), ("lost", 0.5))),      # LOST-heavy: many End/Start pairs
    (20000, 1 << 20, 2, (("switch", 1.0),)),                     # uniform 32-byte records
    (500, 1 << 15, 4, (("comm", 0.4), ("fork", 0.3), ("exit", 0.3))),
])
def test_gpu_switch_events(ctx, n_records, ds, n_cpus, mix):
    d = hbt.switch_desc(tid_level=4)
    rings = []
    for c in range(n_cpus):
        # (fresh FORK tids per CPU: the reference's generators share one thread map)
        s = tg.ctx_switch_stream(tg.stream_seed(40, c) + n_records, n_records, mix=mix,
                                 fork_tid_base=c << 22)
        assert s.numel() * 8 < ds
        tail = tg.random_tail(900 + c, ds)
        ring, head = ring_np(s, ds, tail, fill=5)
        rings.append((ring, tail, head))
    res, ev = gpu_switch(ctx, rings, d, ds, cpu_base=16)
    # with oracle/_ref on the box this runs the REAL CpuThreadSwitchGenerator over the rings
    want = orc.switch_events([r for r, _, _ in rings], [t for _, t, _ in rings],
                             [h for _, _, h in rings], d, cpu_base=16)
    for c in range(n_cpus):
        check(res, ev, want[c], c)
    if "ref_events" in want[0]:
        mapped = orc.map_tids_to_vids(ev)   # the host-side tid -> vid step, applied to the GPU stream
        for c in range(n_cpus):
            assert same_events(mapped[c], want[c]["ref_events"]), f"cpu {c}: GPU events != reference's"
    assert ctx.reruns == 0
    # device view: all CPUs back to back
    base, off = ctx.switch_events_dev()
    assert base and [int(off[c + 1] - off[c]) for c in range(n_cpus)] == [len(e) for e in ev]


@pytest.mark.gpu
def test_gpu_switch_events_reference_stream(ctx):
    """The stream the reference generator itself was driven over (when oracle/_ref is there,
    the GPU events mapped tid -> vid equal the reference's own)."""
    ds = 1 << 16
    rings, tails, heads = consistent_rings(1, 3, 700, ds, 57)
    d = hbt.switch_desc(tid_level=2)
    res, ev = gpu_switch(ctx, [(rings[c], tails[c], heads[c]) for c in range(3)], d, ds, cpu_base=4)
    for c in range(3):
        check(res, ev, orc.port_switch_events(rings[c], tails[c], heads[c], d, cpu=4 + c), c)
    if orc.have_ref():
        ref = orc.ref_switch_events(rings, tails, heads, tid_level=2, cpu_base=4)
        mapped = orc.map_tids_to_vids(ev)
        for c in range(3):
            assert same_events(ref[c]["events"], mapped[c]), c


@pytest.mark.gpu
def test_gpu_switch_events_state_across_calls_and_max_records(ctx):
    ds = 1 << 18
    d = hbt.switch_desc(tid_level=1)
    n_cpus = 3
    streams = [tg.ctx_switch_stream(7000 + c, 4000, mix=(("switch", 0.7), ("lost", 0.2), ("fork", 0.1)))
               for c in range(n_cpus)]
    tails = [tg.random_tail(70 + c, ds) for c in range(n_cpus)]
    full = [ring_np(s, ds, t, fill=9) for s, t in zip(streams, tails)]
    # first call stops after 1500 records, the second drains the rest: state carries over
    d1 = hbt.switch_desc(tid_level=1, max_records=1500)
    rings = [(full[c][0], tails[c], full[c][1]) for c in range(n_cpus)]
    res1, ev1 = gpu_switch(ctx, rings, d1, ds)
    want1 = [orc.port_switch_events(r, t, h, d1, cpu=c) for c, (r, t, h) in enumerate(rings)]
    for c in range(n_cpus):
        check(res1, ev1, want1[c], c)
        assert int(res1[c]["ret"]) == 1500
    rings2 = [(full[c][0], want1[c]["new_tail"], full[c][1]) for c in range(n_cpus)]
    res2, ev2 = gpu_switch(ctx, rings2, d, ds, reset=False)
    for c in range(n_cpus):
        want2 = orc.port_switch_events(full[c][0], want1[c]["new_tail"], full[c][1], d, cpu=c,
                                       last_tstamp=want1[c]["carry_tstamp"],
                                       in_write_errors=want1[c]["num_unknown"],
                                       counters=want1[c]["counters"])
        check(res2, ev2, want2, c)
        # both calls together = one drain
        whole = orc.port_switch_events(full[c][0], tails[c], full[c][1], d, cpu=c)
        assert same_events(np.concatenate([ev1[c], ev2[c]]), whole["events"])


@pytest.mark.gpu
def test_gpu_switch_events_short_record_and_foreign_types(ctx):
    ds = 1 << 19
    d = hbt.switch_desc()
    # SAMPLE / MMAP records in the stream are not handled in this mode: skipped by size
    s = tg.mixed_stream(333, 3000, 4).numpy().view(np.uint64).copy()
    ring, head = ring_np(__import__("torch").from_numpy(s.view(np.int64)), ds, 8 * 31)
    res, ev = gpu_switch(ctx, [(ring, 8 * 31, head)], d, ds)
    check(res, ev, orc.port_switch_events(ring, 8 * 31, head, d), 0)
    assert int(res[0]["ret"]) == -61 and len(ev[0]) > 500
    # a handled record shorter than its struct: -EPERM in front of it, events before it kept
    s2 = tg.ctx_switch_stream(91, 2000, mix=(("switch", 1.0),)).numpy().view(np.uint64).copy()
    k = 4 * 1234
    s2[k] = (s2[k] & np.uint64(0x0000FFFFFFFFFFFF)) | (np.uint64(24) << np.uint64(48))  # 24-byte SWITCH
    ring, head = ring_np(__import__("torch").from_numpy(s2.view(np.int64)), ds, 0)
    res, ev = gpu_switch(ctx, [(ring, 0, head)], d, ds)
    want = orc.port_switch_events(ring, 0, head, d)
    assert want["ret"] == -1 and want["n_records"] == 1234
    check(res, ev, want, 0)


@pytest.mark.gpu
@pytest.mark.parametrize("n_cpus,n_records", [(1, 500), (2, 800), (5, 700), (8, 3000), (13, 1500), (64, 400)])
def test_gpu_switch_events_merged(ctx, n_cpus, n_records):
    """k-way merge of the per-CPU streams = the Combinator tree's order."""
    ds = 1 << 18
    d = hbt.switch_desc(tid_level=1)
    rings = []
    for c in range(n_cpus):
        # no LOST: a WriteErrorsEnd carries the LOST record's own (later) timestamp, all
        # other events are in timestamp order; CPUs share the time base so streams interleave
        n = n_records if c != 1 else 0
        s = tg.ctx_switch_stream(5000 + c, max(n, 1), mix=(("switch", 0.8), ("fork", 0.1), ("exit", 0.1)),
                                 t0=1_000_000 + 37 * c)
        if n == 0:
            s = s[:0]
        tail = tg.random_tail(300 + c, ds)
        ring, head = ring_np(s, ds, tail, fill=3)
        rings.append((ring, tail, head))
    res, ev = gpu_switch(ctx, rings, d, ds, cpu_base=3)
    merged = ctx.fetch_switch_events_merged()
    assert len(merged) == sum(len(e) for e in ev)
    if n_cpus <= 13:
        assert same_events(merged, orc.combinator_merge(ev))
    allev = np.concatenate(ev)
    assert same_events(merged, allev[np.lexsort((allev["comp_unit_id"], allev["tstamp"]))])
    # asking again returns the cached result
    assert same_events(ctx.fetch_switch_events_merged(), merged)


@pytest.mark.gpu
@pytest.mark.parametrize("ds", [4096, 16384, 32768])
def test_gpu_switch_events_small_rings_and_empty_windows(ctx, ds):
    """Rings too small to tile take the one-thread path (k_decode_small); empty windows
    produce no events and leave the generator state alone."""
    d = hbt.switch_desc(tid_level=3)
    rings = []
    for c in range(4):
        n = 0 if c == 2 else (ds // 40) - 8
        s = tg.ctx_switch_stream(600 + c, max(n, 1), mix=(("switch", 0.6), ("lost", 0.2), ("comm", 0.2)))
        s = s[:0] if n == 0 else s[: (ds // 8) - 8]
        # cut at a record boundary: walk the headers
        w = s.numpy().view(np.uint64)
        pos = 0
        while pos < len(w) and pos + int(w[pos] >> np.uint64(51)) <= len(w):
            pos += int(w[pos] >> np.uint64(51))
        s = s[:pos]
        tail = 8 * (ds // 16 + c)
        ring, head = ring_np(s, ds, tail, fill=2)
        rings.append((ring, tail, head))
    res, ev = gpu_switch(ctx, rings, d, ds)
    for c, (ring, tail, head) in enumerate(rings):
        check(res, ev, orc.port_switch_events(ring, tail, head, d, cpu=c), c)
    assert int(res[2]["n_rows"]) == 0 and int(res[2]["ret"]) == -61


@pytest.mark.gpu
def test_gpu_switch_events_payload_like_headers(ctx):
    """COMM names filled with bytes that look like record headers: segment-entry guesses go
    wrong and the batch is re-run with forced entries; the events must not change."""
    ds = 1 << 20
    d = hbt.switch_desc(tid_level=0)
    s = tg.ctx_switch_stream(4242, 20000, mix=(("comm", 0.5), ("switch", 0.5))).numpy().view(np.uint64).copy()
    pos = 0
    fake = np.uint64(15 | (32 << 48))  # a plausible SWITCH_CPU_WIDE header of 32 bytes
    while pos < len(s):
        words = int(s[pos] >> np.uint64(51))
        if int(s[pos] & np.uint64(0xFFFFFFFF)) == 3:  # COMM: hdr, pid/tid, name..., sample_id(2)
            for k in range(2, words - 2):
                s[pos + k] = fake
        pos += words
    import torch
    ring, head = ring_np(torch.from_numpy(s.view(np.int64)), ds, 8 * 5)
    res, ev = gpu_switch(ctx, [(ring, 8 * 5, head)], d, ds)
    check(res, ev, orc.port_switch_events(ring, 8 * 5, head, d), 0)
