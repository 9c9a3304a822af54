"""Slices downstream of the decode: the interval join, the per-stack-id slice
frequencies and the interval x bin duration matrix on the GPU (hbt_slices_join_dev,
hbt_slice_stats_*, hbt_slice_bins_*).

hbt/src/tagstack needs Boost.Intrusive and MonData.h the absent hbt/src/phase, so the reference's own
code for this row does not compile here.  The checker is the plain-Python restatement in
tests/_oracle.py; for the join it is pinned to the exact expectations of the reference's own
hbt/src/tagstack/tests/IntervalSlicerTest.cpp (below, test by test), and the CUDA path is held to those
expectations directly as well.  stackIdFreqs and TagStackIdBinner have no reference test: parity for
them rests on the restatement (unpinned)."""
import numpy as np
import pytest

import dynolog_1_b200 as hbt
import _oracle as orc

A, TP, TY, PC = hbt.TT_ANALYSIS, hbt.TT_THREAD_PREEMPTED, hbt.TT_THREAD_YIELD, hbt.TT_PHASE_CHANGE

# (slices written so far, offset, tstamp, duration, trim, abort) -> (expected slices or None, consumed):
# hbt/src/tagstack/tests/IntervalSlicerTest.cpp:13-200 (makeIntervalIntersectingSlices) and :202-430
# (consumeSlices; expected consumer marks as offset + consumed)
S1 = (10, 20, 1, A, A)
S1B = (0, 99, 1, A, A)
S2B = (99, 2, 1, A, A)
M1, M2, M3, M4 = (0, 99, 1, A, A), (99, 2, 33, TP, PC), (101, 1, 1, A, A), (200, 10, 1, PC, PC)
KNOWN = [
    ([], 0, 0, 100, False, False, None, 0),                                   # EmptySlices
    ([S1], 0, 0, 100, False, False, [S1], 1),                                 # SingleConsumerOneSliceNoTrim
    ([S1], 0, 0, 100, True, True, None, 0),                                   # ...OneSliceTrim
    ([(10, 90, 1, A, A)], 0, 0, 100, True, True, None, 0),                    # ...MissingStart
    ([S1B], 0, 0, 100, True, True, None, 0),                                  # ...MissingEnd
    ([S1B], 0, 0, 100, True, False, [S1B], 1),                                # ...MissingEndNoFail
    ([S1B, S2B], 0, 0, 100, True, True, [S1B, (99, 1, 1, A, A)], 1),          # SingleConsumerMultiple
    ([S1B, S2B], 1, 0, 101, True, True, None, 0),
    ([S1B, S2B], 0, 99, 101, True, True, None, 0),
    ([S1B, S2B], 0, 100, 102, True, True, None, 0),
    ([S1B, S2B], 0, 100, 1, True, True, [(100, 1, 1, A, A)], 2),
    ([M1], 0, 0, 100, True, True, None, 0),                                   # MultipleSlicerInSequence
    ([M1, M2], 0, 0, 99, True, True, [M1], 1),
    ([M1, M2], 1, 0, 101, True, True, None, 0),
    ([M1, M2], 1, 99, 101, True, True, None, 0),
    ([M1, M2], 1, 100, 102, True, True, None, 0),
    ([M1, M2], 1, 100, 1, True, True, [(100, 1, 33, A, PC)], 1),
    ([M1, M2, M3], 1, 100, 1, True, True, [(100, 1, 33, A, PC)], 1),
    ([M1, M2, M3, M4], 2, 100, 200, True, True, None, 0),
    ([M1, M2, M3, M4], 2, 100, 200, True, False, [M3, M4], 2),
    ([M1, M2, M3, M4], 2, 100, 105, True, False, [M3, (200, 5, 1, PC, A)], 1),
]


def to_np(slices):
    a = np.zeros(len(slices), dtype=hbt.SLICE_DTYPE)
    for i, s in enumerate(slices):
        a[i]["tstamp"], a[i]["duration"], a[i]["stack_id"], a[i]["swin_type"], a[i]["swout_type"] = s
    return a


def from_np(a):
    return [(int(x["tstamp"]), int(x["duration"]), int(x["stack_id"]), int(x["swin_type"]), int(x["swout_type"]))
            for x in a]


@pytest.mark.parametrize("case", range(len(KNOWN)))
def test_restatement_reproduces_interval_slicer_test(case):
    slices, off, t, d, trim, abort, want, consumed = KNOWN[case]
    got, n = orc.port_interval_slices(slices, off, t, d, trim, abort)
    assert got == want and n == consumed


def random_slices(seed, n, n_ids, gap_every=0):
    rng = np.random.default_rng(seed)
    t = 1000
    out = []
    for i in range(n):
        d = int(rng.integers(1, 400))
        if gap_every and i % gap_every == gap_every - 1:
            t += int(rng.integers(1, 300))          # a hole in the coverage
        out.append((t, d, int(rng.integers(0, n_ids)) if rng.random() > 0.01 else 2**64 - 1,
                    int(rng.integers(1, 5)), int(rng.integers(1, 5))))
        t += d
    return out


def intervals_over(slices, seed, n, aligned=False):
    rng = np.random.default_rng(seed)
    t0, t1 = slices[0][0], slices[-1][0] + slices[-1][1]
    step = max(1, (t1 - t0) // n)
    out, t = [], t0 - (0 if aligned else 50)
    for _ in range(n + 3):
        d = step if aligned else int(rng.integers(1, 2 * step))
        out.append((t, d))
        t += d if rng.random() > 0.1 else d + int(rng.integers(0, step))   # count intervals may leave gaps
    return out


def gpu_join(st, slices, intervals, trim, abort, offset=0, stride=2):
    import torch
    sl = torch.from_numpy(to_np(slices).view(np.uint8)).cuda() if slices else torch.zeros(32, dtype=torch.uint8, device="cuda")
    iv = np.zeros((len(intervals), stride), dtype=np.uint64)
    for k, (t, d) in enumerate(intervals):
        iv[k, 0], iv[k, 1] = t, d
    ivd = torch.from_numpy(iv.view(np.int64)).cuda()
    cap = len(slices) + 2 * len(intervals) + 1
    out = torch.zeros(cap * 32, dtype=torch.uint8, device="cuda")
    offs = torch.zeros(len(intervals) + 1, dtype=torch.int64, device="cuda")
    n, mark = st.join_dev(sl, len(slices), offset, ivd, stride, len(intervals), trim, abort, out, cap, offs)
    st.ctx.sync()
    o = offs.cpu().numpy().astype(np.int64)
    flat = out.cpu().numpy()[:n * 32].view(hbt.SLICE_DTYPE)
    return [from_np(flat[o[k]:o[k + 1]]) for k in range(len(intervals))], mark, (out, offs, n)


@pytest.mark.gpu
def test_gpu_join_reproduces_interval_slicer_test(built):
    ctx = hbt.Context(0)
    st = hbt.SliceStats(ctx, 64)
    try:
        for slices, off, t, d, trim, abort, want, consumed in KNOWN:
            got, mark, _ = gpu_join(st, slices, [(t, d)], trim, abort, offset=off)
            assert got[0] == (want or []), (slices, off, t, d, trim, abort)
            assert mark == off + consumed
    finally:
        st.close()
        ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("trim,abort", [(True, True), (True, False), (False, False), (False, True)])
@pytest.mark.parametrize("seed,n,gap_every,aligned", [(1, 3000, 0, False), (2, 3000, 37, False), (3, 500, 0, True),
                                                      (4, 20000, 501, False)])
def test_gpu_join_random(built, trim, abort, seed, n, gap_every, aligned):
    slices = random_slices(seed, n, 50, gap_every)
    intervals = intervals_over(slices, seed + 100, max(8, n // 12), aligned)
    import errno
    try:
        want, want_mark = orc.port_consume_slices(slices, intervals, trim, abort)
    except orc.ReferenceThrows:
        want = None   # untrimmed slices reaching across an interval's start: the reference throws
    assert trim is False or want is not None   # with trimming the slices never outgrow the interval
    ctx = hbt.Context(0)
    st = hbt.SliceStats(ctx, 64)
    try:
        if want is None:
            with pytest.raises(hbt.HbtError) as e:
                gpu_join(st, slices, intervals, trim, abort, stride=5)
            assert e.value.code == -errno.EDOM
            # with holes between the slices nothing reaches across and the untrimmed join is defined
            slices = [(t, max(1, d // 2), sid, a, b) for (t, d, sid, a, b) in slices]
            intervals = [(slices[i][0], slices[min(i + 7, len(slices) - 1)][0] - slices[i][0] + 3)
                         for i in range(0, len(slices) - 8, 8)]
            want, want_mark = orc.port_consume_slices(slices, intervals, trim, abort)
            assert abort or sum(len(w) for w in want) > len(intervals)   # (holes abort every interval)
        got, mark, _ = gpu_join(st, slices, intervals, trim, abort, stride=5)   # intervals inside wider rows
        assert got == want
        assert mark == want_mark
    finally:
        st.close()
        ctx.close()


@pytest.mark.gpu
def test_gpu_freqs_and_bin_matrix(built):
    """Three compute units: join, stackIdFreqs per unit accumulated over the units, then the
    TagStackIdBinner matrix (bins compared through their stack ids)."""
    ctx = hbt.Context(0)
    st = hbt.SliceStats(ctx, 4096)
    try:
        units, keep = [], []
        n_iv = 0
        for u in range(3):
            slices = random_slices(50 + u, 4000, 300 if u else 40, gap_every=0 if u != 1 else 97)
            # one unit keeps zero-duration trims out of reach, one has intervals beyond its slices
            intervals = intervals_over(slices, 70 + u, 120 + 10 * u)
            want, _ = orc.port_consume_slices(slices, intervals, True, False)
            got, _, bufs = gpu_join(st, slices, intervals, True, False)
            assert got == want
            st.add_unit_dev(bufs[0], bufs[1], len(intervals), bufs[2])
            units.append(want)
            keep.append((bufs, len(intervals)))
            n_iv = max(n_iv, len(intervals))
        want_f = orc.port_accum_freqs([orc.port_stack_id_freqs(w) for w in units])
        got_f = st.fetch()
        assert [int(x) for x in got_f["stack_id"]] == sorted(want_f)
        for x in got_f:
            assert [int(x["duration"]), int(x["num_obs"]), int(x["num_intervals"])] == want_f[int(x["stack_id"])]
        assert 2**64 - 1 in want_f   # kInvalidTagStackId is a key like any other
        # bin matrix
        min_iv, min_dur = 5, 2000
        n_bins, sids, bids = st.bins_prepare(min_iv, min_dur, n_iv)
        for bufs, n in keep:
            st.bins_add_unit_dev(bufs[0], bufs[1], n)
        m = st.bins_fetch()
        cols = orc.port_bin_matrix(units, want_f, min_iv, min_dur, n_iv)
        assert n_bins == len(cols)               # one bin per kept stack id + the ignored bin 0
        assert sorted(int(b) for b in bids if b) == list(range(1, n_bins))
        for sid, b in zip(sids, bids):
            if b:
                assert [int(v) for v in m[:, b]] == cols[int(sid)], int(sid)
            else:
                assert int(sid) not in cols
        assert [int(v) for v in m[:, 0]] == cols[None]
    finally:
        st.close()
        ctx.close()


@pytest.mark.gpu
def test_gpu_slices_errors(built):
    import errno
    import torch
    ctx = hbt.Context(0)
    st = hbt.SliceStats(ctx, 4)   # far too few stack ids for what follows
    try:
        slices = random_slices(9, 2000, 500)
        intervals = intervals_over(slices, 10, 20)
        _, _, bufs = gpu_join(st, slices, intervals, True, False)
        with pytest.raises(hbt.HbtError) as e:
            st.add_unit_dev(bufs[0], bufs[1], len(intervals), bufs[2])
        assert e.value.code == -errno.ENOSPC
        with pytest.raises(hbt.HbtError) as e:   # an interval of zero duration: HBT_ARG_CHECK_GT(duration, 0)
            gpu_join(st, slices, [(slices[0][0], 0)], True, False)
        assert e.value.code == -errno.EINVAL
    finally:
        st.close()
        ctx.close()
