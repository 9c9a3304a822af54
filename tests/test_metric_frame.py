"""Metric tables: dynolog's MetricFrame / MetricSeries window statistics on the
GPU (hbt_frame_*).  Oracle = the reference's own classes (dynolog/src/metric_frame/*.cpp compiled
into oracle/_ref); the numpy restatement in tests/_oracle.py is pinned to them here, to the
reference's own test expectations (dynolog/tests/metric_frame/MetricFrameTest.cpp:39-167,
MetricFrameTsUnitTest.cpp) and to tests/golden/metric_frame.npz (generated from the reference by
tests/golden/make_golden_frame.py) for machines without the reference tree."""
import os

import numpy as np
import pytest

import dynolog_1_b200 as hbt
import _oracle as orc

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "metric_frame.npz")
S60 = 60 * 10**9


def smoke_frame():
    """The frame of MetricFrameTest.cpp: interval 60 s, length 10, an int64 and a double series;
    sample i = (42 + i, 23.9f + i) at now + 60 s * i for i = 0..10 (the first one falls out)."""
    now = 5_000 * 10**9
    rows = np.zeros((11, 2), dtype=np.uint64)
    rows[:, 0] = np.arange(42, 53, dtype=np.int64).view(np.uint64)
    rows[:, 1] = (np.float32(23.9).astype(np.float64) + np.arange(11)).view(np.uint64)
    return now, rows


def random_case(seed, n_rows, S, frame_len, interval_us, n_q):
    rng = np.random.default_rng(seed)
    is_double = rng.integers(0, 2, S).astype(np.uint8)
    rows = np.zeros((n_rows, S), dtype=np.uint64)
    for s in range(S):
        if is_double[s]:
            rows[:, s] = np.cumsum(rng.random(n_rows) * 1e6 - 2e5).view(np.uint64)
        else:
            rows[:, s] = np.cumsum(rng.integers(-10**6, 10**9, n_rows)).astype(np.int64).view(np.uint64)
    t_last = 10**15 + int(rng.integers(0, 10**9))
    held = min(n_rows, frame_len)
    span = held * interval_us * 1000
    queries = []
    for _ in range(n_q):
        a = t_last - int(rng.integers(-span // 4, span + span // 4))
        b = a + int(rng.integers(0, span))
        queries.append((a, b, int(rng.integers(0, 3)), int(rng.integers(0, 3))))
    queries.append((t_last, t_last, 0, 0))                       # a one-sample window
    queries.append((t_last - span * 3, t_last + span * 3, 0, 0))  # clamps to the whole frame
    return rows, is_double, t_last, queries


def same_bits(a, b):
    return np.array_equal(np.ascontiguousarray(a).view(np.uint64), np.ascontiguousarray(b).view(np.uint64))


def test_restatement_reproduces_the_reference_tests_expectations(built):
    now, rows = smoke_frame()
    length, ranges, o = orc.port_frame_query(rows, [0, 1], 10, 60_000_000, now + 10 * S60,
                                             [(now + 100 * 10**9, now + 400 * 10**9, 0, 0)], 10**9, 0.2)
    assert length == 10 and ranges[0][:2] == (1, 6)            # 120 s .. 420 s
    assert ranges[0][2:] == (now + 2 * S60, now + 7 * S60)
    assert int(o["avg_i64"][0, 0]) == 46 and abs(o["avg_f64"][0, 0] - 46.5) < 1e-3
    assert int(o["pct"][0, 0]) == 45 and int(o["diff"][0, 0]) == 5
    assert abs(o["rate_f64"][0, 0] - 0.01667) < 1e-3
    _, _, o1 = orc.port_frame_query(rows, [0, 1], 10, 60_000_000, now + 10 * S60,
                                    [(now + 100 * 10**9, now + 400 * 10**9, 0, 0)], 60 * 10**9, 0.4)
    assert int(o1["rate_i64"][0, 0]) == 1 and int(o1["rate_i64"][0, 1]) == 1
    assert abs(o1["avg_f64"][0, 1] - 28.4) < 1e-3
    assert abs(o1["pct"][0, 1:].view(np.float64)[0] - 27.9) < 1e-3
    assert abs(o1["diff"][0, 1:].view(np.float64)[0] - 5.0) < 1e-3
    # MetricFrameTsUnitTest: before the first sample PREV_CLOSEST finds nothing, NEXT_CLOSEST the first
    assert orc.port_frame_range(10, now + 10 * S60, 60_000_000, now, now + 140 * 10**9, 1, 0) is None
    assert orc.port_frame_range(10, now + 10 * S60, 60_000_000, now, now + 140 * 10**9, 2, 0)[:2] == (0, 1)


@pytest.mark.parametrize("seed,n_rows,S,frame_len,interval_us", [(1, 50, 3, 64, 1_000_000), (2, 300, 5, 100, 250_000),
                                                                (3, 7, 2, 10, 60_000_000), (4, 1000, 4, 333, 1000)])
def test_restatement_equals_reference(built, seed, n_rows, S, frame_len, interval_us):
    if not orc.have_ref():
        pytest.skip("oracle/_ref not built (reference tree absent)")
    rows, isd, t_last, queries = random_case(seed, n_rows, S, frame_len, interval_us, 40)
    orc.frame_query(rows, isd, frame_len, interval_us, t_last, queries, 10**9, 0.57)   # asserts inside
    orc.frame_query(rows, isd, frame_len, interval_us, t_last, queries, 7 * 10**6, 0.0)


def test_restatement_reproduces_golden(built):
    g = np.load(GOLDEN, allow_pickle=False)
    queries = [tuple(int(x) for x in q) for q in g["queries"]]
    length, ranges, o = orc.port_frame_query(g["rows"], g["is_double"], int(g["frame_len"]), int(g["interval_us"]),
                                             int(g["t_last_ns"]), queries, int(g["period_ns"]), float(g["percentage"]))
    assert length == int(g["length"])
    live = g["ok"].astype(bool)
    assert [r is not None for r in ranges] == list(live)
    for i, r in enumerate(ranges):
        if r is not None:
            assert r == tuple(int(x) for x in g["ranges"][i])
    for k in ("diff", "rate_f64", "rate_i64", "avg_f64", "avg_i64", "pct"):
        assert same_bits(o[k][live], g[k][live]), k


@pytest.mark.gpu
@pytest.mark.parametrize("seed,n_rows,S,frame_len,interval_us", [(11, 50, 3, 64, 1_000_000), (12, 5000, 6, 1024, 250_000),
                                                                (13, 7, 2, 10, 60_000_000), (14, 20000, 9, 8192, 1000),
                                                                (15, 1, 1, 4, 5_000)])
def test_gpu_frame_against_the_reference(built, seed, n_rows, S, frame_len, interval_us):
    rows, isd, t_last, queries = random_case(seed, n_rows, S, frame_len, interval_us, 60)
    for period_ns, pct in ((10**9, 0.57), (3 * 10**6, 1.0)):
        length, want_ranges, want = orc.frame_query(rows, isd, frame_len, interval_us, t_last, queries, period_ns, pct)
        ctx = hbt.Context(0)
        fr = hbt.Frame(ctx, isd, frame_len, interval_us)
        try:
            # in two batches: the ring wraps across calls exactly like sample-by-sample addSample
            cut = n_rows // 3
            if cut:
                fr.add_samples(rows[:cut], t_last - (n_rows - cut) * interval_us * 1000)
            fr.add_samples(rows[cut:], t_last)
            assert len(fr) == length
            got_ranges = [fr.get_range(*q) for q in queries]
            assert got_ranges == want_ranges
            live = [i for i, r in enumerate(want_ranges) if r is not None]
            got = fr.query([got_ranges[i] for i in live], period_ns, pct)
            for k in want:
                assert same_bits(got[k], want[k][live]), k   # integers and doubles: bit-exact
        finally:
            fr.close()
            ctx.close()


@pytest.mark.gpu
def test_gpu_frame_fed_from_the_counts_pipeline(built):
    """The tables of the counter-group kernels go into a frame without leaving the GPU: scaled counts
    (int64) of every interval appended with hbt_frame_add_samples_dev, then window diff / rate."""
    import torch
    from dynolog_1_b200 import tracegen as tg
    I, Cc, E = 500, 16, 4
    reads = tg.counts_trace(77, I, Cc, E, device="cuda", multiplexed=True)
    scaled = torch.zeros((I, E), dtype=torch.int64, device="cuda")
    ctx = hbt.Context(0)
    fr = hbt.Frame(ctx, [0] * E, 256, 1_000_000)
    try:
        ctx.counts_pipeline_dev(reads, I, Cc, E, [], scaled=scaled)
        t_last = 10**12
        fr.add_samples_dev(scaled, I, E, t_last)
        ctx.sync()
        want_scaled = orc.counts(reads.cpu().numpy().view(np.uint64))["scaled"]
        queries = [(t_last - 100 * 10**9, t_last, 0, 0), (t_last - 10**15, t_last - 200 * 10**9, 0, 0)]
        _, want_ranges, want = orc.frame_query(want_scaled, [0] * E, 256, 1_000_000, t_last, queries, 10**9, 0.5)
        ranges = [fr.get_range(*q) for q in queries]
        assert ranges == want_ranges
        got = fr.query(ranges, 10**9, 0.5)
        for k in want:
            assert same_bits(got[k], want[k]), k
    finally:
        fr.close()
        ctx.close()


@pytest.mark.gpu
def test_gpu_frame_argument_errors(built):
    import errno
    ctx = hbt.Context(0)
    try:
        with pytest.raises(hbt.HbtError) as e:
            hbt.Frame(ctx, [0], 16, 0)           # interval 0: the reference's constructor throws
        assert e.value.code == -errno.EINVAL
        with pytest.raises(hbt.HbtError):
            hbt.Frame(ctx, [0], hbt.FRAME_MAX_LEN + 1, 1000)
        fr = hbt.Frame(ctx, [0, 1], 8, 1000)
        assert fr.get_range(0, 10, 0, 0) is None  # empty frame: std::nullopt
        fr.add_samples(np.arange(6, dtype=np.uint64).reshape(3, 2), 10**9)
        with pytest.raises(hbt.HbtError):
            fr.query([(0, 5, 0, 0)], 10**9, 0.5)  # window outside the 3 samples held
        fr.close()
    finally:
        ctx.close()
