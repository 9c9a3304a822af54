"""Pins the C restatement (oracle/hbt_oracle.c) to the REFERENCE's own code
(oracle/_ref/libhbt_ref.so = reference headers driven by oracle/ref_harness.cpp)
on seeded inputs.  Skipped where oracle/_ref could not be built or shipped;
tests/test_golden.py covers the same ground from committed vectors."""
import numpy as np
import pytest

import dynolog_1_b200 as hbt
from dynolog_1_b200 import tracegen as tg
import _oracle as orc

pytestmark = pytest.mark.skipif(not orc.have_ref(), reason="oracle/_ref not built")


def ring_np(stream, ds, tail, fill=77):
    ring, head = tg.place_in_ring(stream, ds, tail, fill_seed=fill)
    return ring.numpy().view(np.uint64), head


def assert_same(a, b, keys=("ret", "n_records", "new_tail", "n_rows", "counters", "num_unknown",
                            "carry_tstamp")):
    for k in keys:
        assert a[k] == b[k], (k, a[k], b[k])
    if "rows" in a and "rows" in b:
        assert np.array_equal(a["rows"], b["rows"])
    if "carry_last" in a and "carry_last" in b:
        assert np.array_equal(a["carry_last"], b["carry_last"])
    if a.get("hist") is not None and b.get("hist") is not None:
        assert a["hist"] == b["hist"]


@pytest.mark.parametrize("n_events,lost_every,max_records", [
    (8, 0, hbt.DRAIN), (8, 50, hbt.DRAIN), (1, 7, hbt.DRAIN), (32, 13, hbt.DRAIN),
    (64, 0, hbt.DRAIN), (8, 50, 100), (8, 50, 1), (8, 0, 0)])
def test_sample_deltas(built, n_events, lost_every, max_records):
    ds = 1 << 17
    n = (ds - 4096) // (40 + 8 * n_events + 3)
    s = tg.group_sample_stream(tg.stream_seed(3, n_events), n, n_events, lost_every=lost_every)
    tail = tg.random_tail(n_events + lost_every, ds)
    ring, head = ring_np(s, ds, tail)
    d = hbt.sampling_desc(n_events, max_records=max_records)
    a = orc.ref_consume(ring, tail, head, d)
    b = orc.port_consume(ring, tail, head, d)
    assert_same(a, b)
    if max_records == hbt.DRAIN:
        assert a["ret"] == -61  # -ENODATA: a full drain never returns the count
        assert a["counters"][hbt.CNT_SAMPLE] == n


def test_sample_deltas_carry_across_calls(built):
    """Two consume() calls over consecutive windows == one call over both."""
    ds, E = 1 << 16, 4
    s = tg.group_sample_stream(11, 500, E, lost_every=40)
    tail = tg.random_tail(5, ds)
    ring, head = ring_np(s, ds, tail)
    d = hbt.sampling_desc(E, max_records=123)
    for impl in (orc.ref_consume, orc.port_consume):
        r1 = impl(ring, tail, head, d, carry_tstamp=12345, carry_last=np.arange(E, dtype=np.uint64))
        d2 = hbt.sampling_desc(E)
        r2 = impl(ring, r1["new_tail"], head, d2, carry_tstamp=r1["carry_tstamp"],
                  carry_last=r1["carry_last"], counters=r1["counters"],
                  num_unknown=r1["num_unknown"])
        whole = impl(ring, tail, head, d2, carry_tstamp=12345,
                     carry_last=np.arange(E, dtype=np.uint64))
        assert r1["ret"] == 123
        assert np.array_equal(np.concatenate([r1["rows"], r2["rows"]]), whole["rows"])
        assert r2["counters"] == whole["counters"] and r2["num_unknown"] == whole["num_unknown"]
    assert_same(orc.ref_consume(ring, tail, head, d), orc.port_consume(ring, tail, head, d))


def test_timestamp_regression_and_bad_size(built):
    ds, E = 1 << 16, 8
    s = tg.group_sample_stream(21, 300, E, regress_at=200)
    ring, head = ring_np(s, ds, 8 * 1000)
    d = hbt.sampling_desc(E)
    a, b = orc.ref_consume(ring, 8000, head, d), orc.port_consume(ring, 8000, head, d)
    assert a["ret"] == -1 and a["n_records"] == 200 and a["carry_tstamp"] == hbt.INVALID_TS
    assert_same(a, b)
    # SAMPLE whose header.size disagrees with nr: -EPERM before the handler
    s2 = tg.group_sample_stream(22, 300, E)
    s2[(5 + E) * 150 + 2] = E + 1
    ring, head = ring_np(s2, ds, 0)
    a, b = orc.ref_consume(ring, 0, head, d), orc.port_consume(ring, 0, head, d)
    assert a["ret"] == -1 and a["n_records"] == 150 and a["carry_tstamp"] != hbt.INVALID_TS
    assert_same(a, b)


def test_tsc_transform(built):
    ds, E = 1 << 16, 6
    s = tg.group_sample_stream(31, 400, E, lost_every=90)
    ring, head = ring_np(s, ds, 4096 * 3 + 64)
    d = hbt.sampling_desc(E, tsc_mask=0b100101, tsc=(987654321, 1234567, 21))
    assert_same(orc.ref_consume(ring, 4096 * 3 + 64, head, d),
                orc.port_consume(ring, 4096 * 3 + 64, head, d))
    for v in (0, 1, 2**40 + 12345, 2**63 + 99, 2**64 - 1):
        assert orc.ref().hbtref_kernel_time_from_tsc(v, d.tsc) == \
            orc.port().hbt_oracle_kernel_time_from_tsc(v, d.tsc)


@pytest.mark.parametrize("mode", [hbt.MODE_SAMPLING, hbt.MODE_CONTEXT_SWITCH, hbt.MODE_AUX_SPACE,
                                  hbt.MODE_MIXED])
@pytest.mark.parametrize("max_records", [hbt.DRAIN, 777])
def test_mixed_stream_counters(built, mode, max_records):
    """Mixed SAMPLE/SWITCH/COMM/MMAP/LOST/FORK/EXIT stream through every handler
    set: which counters move depends on the mode's non-void record types."""
    ds, E = 1 << 18, 8
    s = tg.mixed_stream(41, 2500, E)
    tail = tg.random_tail(9, ds)
    ring, head = ring_np(s, ds, tail)
    d = hbt.count_desc(mode, max_records=max_records)
    a, b = orc.ref_consume(ring, tail, head, d), orc.port_consume(ring, tail, head, d)
    assert_same(a, b, keys=("ret", "n_records", "new_tail", "counters"))
    if max_records == hbt.DRAIN and mode == hbt.MODE_MIXED:
        assert a["n_records"] == 2500 and a["counters"][hbt.CNT_SWITCH_CPU_WIDE] > 0


def test_mixed_stream_deltas(built):
    ds, E = 1 << 18, 8
    s = tg.mixed_stream(43, 2500, E, mix=(("sample", 0.6), ("switch", 0.2), ("comm", 0.05),
                                          ("mmap", 0.05), ("lost", 0.02), ("fork", 0.04),
                                          ("exit", 0.04)))
    tail = tg.random_tail(10, ds)
    ring, head = ring_np(s, ds, tail)
    d = hbt.sampling_desc(E, handled=hbt.MODE_MIXED)
    assert_same(orc.ref_consume(ring, tail, head, d), orc.port_consume(ring, tail, head, d))


@pytest.mark.parametrize("words", [5, 6, 16])
def test_ip_histogram(built, words):
    ds = 1 << 18
    n = (ds - 8192) // (8 * words)
    s = tg.ip_sample_stream(51, n, cpu=3, sample_words=words, n_pids=64, n_ips=256, lost_every=97)
    tail = tg.random_tail(12, ds)
    ring, head = ring_np(s, ds, tail)
    d = hbt.ip_desc(8 * words)
    a, b = orc.ref_consume(ring, tail, head, d, cpu=3), orc.port_consume(ring, tail, head, d, cpu=3)
    assert_same(a, b, keys=("ret", "n_records", "new_tail", "counters"))
    assert sum(a["hist"].values()) == n


@pytest.mark.parametrize("I,C,E,mux", [(50, 64, 8, False), (40, 7, 32, True), (30, 3, 1, True),
                                       (5, 512, 2, True), (20, 16, 64, True)])
def test_counts(built, I, C, E, mux):
    reads = tg.counts_trace(61 + E, I, C, E, multiplexed=mux).numpy().view(np.uint64)
    ratios = [(min(1, E - 1), 0), (E - 1, min(1, E - 1))]
    a, b = orc.ref_counts(reads, ratios), orc.port_counts(reads, ratios)
    for k in ("sums", "scaled", "deltas"):
        assert np.array_equal(a[k], b[k]), k
    assert np.array_equal(a["rates"], b["rates"])
    assert np.array_equal(a["ratios"], b["ratios"])  # same IEEE ops: bit-exact on the host
    if not mux:
        # no multiplexing: scaled ~ raw sum, but NOT always equal -- count*enabled
        # rounds in double once it passes 2^53 (CpuEventsGroup.h:455-458)
        raw = a["sums"][:, 3:].astype(np.float64)
        assert np.all(np.abs(a["scaled"].astype(np.float64) - raw) <= 1e-9 * raw + 2)


def test_counts_edge_values(built):
    """u64 wrap in accum, counts above 2^53, time_enabled == 0, a failed read."""
    I, C, E = 6, 4, 3
    reads = tg.counts_trace(71, I, C, E, multiplexed=True).numpy().view(np.uint64).copy()
    reads[1, :, 3] = np.uint64(2**62)           # sum wraps mod 2^64
    reads[2, :, 4] = np.uint64(2**60 + 12345)   # > 2^53: double rounding matters
    reads[3, :, 1] = 0                          # time_enabled == 0 -> getCount == 0
    reads[3, :, 2] = 0
    reads[4, 2, 0] = 0                          # this CPU's read failed -> interval nullopt
    a, b = orc.ref_counts(reads), orc.port_counts(reads)
    ok = [0, 1, 2, 3, 5]
    assert np.array_equal(a["sums"][ok], b["sums"][ok])
    assert a["sums"][4, 0] == 0 and b["sums"][4, 0] == 0
    assert np.array_equal(a["scaled"], b["scaled"]) and np.array_equal(a["deltas"], b["deltas"])
    assert np.all(a["scaled"][3] == 0)
    threads = orc.ref_counts(reads, threads=3)
    assert np.array_equal(threads["scaled"], a["scaled"])
    assert np.array_equal(threads["deltas"], a["deltas"])


@pytest.mark.parametrize("budget", [0, 1, 7, 150, 10**6])
def test_full_output_ring_stops_consume_in_front_of_the_sample(built, budget):
    """Back-pressure (PerCpuCountSampleGenerator.h:182-195, CpuEventsGroup.h:1441-1444): the real
    generator with room for `budget` rows against the restatement's rows_cap, then the rest of the
    window once the ring was drained (the refused SAMPLE re-seeds)."""
    if not orc.have_ref():
        pytest.skip("oracle/_ref not built (reference tree absent)")
    ds, E = 1 << 17, 4
    s = tg.group_sample_stream(505, 900, E, lost_every=61)
    tail = 3 * ds + ds - 4096
    ring, head = tg.place_in_ring(s, ds, tail, fill_seed=3)
    ring = ring.numpy().view(np.uint64)
    d = hbt.sampling_desc(E)
    r = orc.ref_consume_budget(ring, tail, head, d, budget, carry_tstamp=1000, carry_last=np.zeros(E, np.uint64))
    p = orc.port_consume(ring, tail, head, d, 1000, np.zeros(E, np.uint64), rows_budget=budget)
    for k in ("ret", "n_records", "new_tail", "counters", "n_rows", "num_unknown", "carry_tstamp"):
        assert r[k] == p[k], k
    assert np.array_equal(r["rows"], p["rows"]) and np.array_equal(r["carry_last"], p["carry_last"])
    if budget < 800:
        assert r["ret"] == -28 and r["n_rows"] == budget and r["carry_tstamp"] == hbt.INVALID_TS
        # drained: the rest of the window
        r2 = orc.ref_consume_budget(ring, r["new_tail"], head, d, 10**6, r["carry_tstamp"], r["carry_last"],
                                    r["counters"], r["num_unknown"])
        p2 = orc.port_consume(ring, p["new_tail"], head, d, p["carry_tstamp"], p["carry_last"], p["counters"],
                              p["num_unknown"])
        for k in ("ret", "n_records", "new_tail", "counters", "n_rows", "num_unknown", "carry_tstamp"):
            assert r2[k] == p2[k], k
        assert np.array_equal(r2["rows"], p2["rows"])
        assert r2["num_unknown"] >= r["num_unknown"] + 1   # the refused SAMPLE only re-seeds
