"""Golden vectors in tests/golden/*.npz were produced by the REFERENCE's own code
(tests/golden/make_golden.py over oracle/_ref).  The oracle port must reproduce
them anywhere (no GPU, no reference tree); the CUDA path must too (-m gpu)."""
import glob
import os

import numpy as np
import pytest

import dynolog_1_b200 as hbt
import _oracle as orc

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
DECODE_CASES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "*.npz"))
                      if not os.path.basename(p).startswith(("counts_", "switch_", "metric_")))
COUNTS_CASES = ["counts_mux", "counts_c1_shape"]
SWITCH_CASES = ["switch_events_lost", "switch_events_plain"]


def load_desc(f):
    v = [int(x) for x in f["desc"]]
    d = hbt.DecodeDesc()
    (d.handled_mask, d.sample_layout, d.aggregate, d.n_events, d.ip_sample_bytes, d.max_records,
     d.tsc_transform_mask) = v[:7]
    d.tsc = hbt.TscParams(v[7], v[8], v[9], 0)
    return d


def want_of(f):
    w = dict(ret=int(f["ret"]), n_records=int(f["n_records"]), new_tail=int(f["new_tail"]),
             n_rows=int(f["n_rows"]), counters=[int(x) for x in f["counters"]],
             num_unknown=int(f["num_unknown"]), carry_tstamp=int(f["carry_tstamp"]))
    if "rows" in f:
        w["rows"] = f["rows"]
        w["carry_last"] = f["carry_last"]
    if "hist_keys" in f:
        w["hist"] = {tuple(int(x) for x in k): int(c) for k, c in zip(f["hist_keys"], f["hist_counts"])}
    return w


def test_golden_files_present():
    assert len(DECODE_CASES) >= 12


@pytest.mark.parametrize("case", DECODE_CASES)
def test_oracle_port_matches_reference_vectors(built, case):
    f = np.load(os.path.join(GOLDEN, case + ".npz"))
    d = load_desc(f)
    tail, head = (int(x) for x in f["window"])
    carry_last = f["carry_in_last"] if "carry_in_last" in f else None
    got = orc.port_consume(f["ring"], tail, head, d, int(f["carry_in_tstamp"]), carry_last,
                           cpu=int(f["cpu"]))
    want = want_of(f)
    keys = ["ret", "n_records", "new_tail", "counters"]
    if d.aggregate == hbt.AGG_COUNT_DELTAS:
        keys += ["n_rows", "num_unknown", "carry_tstamp"]
    for k in keys:
        assert got[k] == want[k], (case, k, got[k], want[k])
    if "rows" in want:
        assert np.array_equal(got["rows"], want["rows"])
        assert np.array_equal(got["carry_last"], want["carry_last"])
    if "hist" in want:
        assert got["hist"] == want["hist"]


@pytest.mark.parametrize("case", COUNTS_CASES)
def test_oracle_counts_match_reference_vectors(built, case):
    f = np.load(os.path.join(GOLDEN, case + ".npz"))
    ratios = [tuple(int(x) for x in r) for r in f["ratios"]]
    got = orc.port_counts(f["reads"], ratios)
    ok = f["sums"][:, 0] != 0  # a failed read leaves the other words unspecified
    assert np.array_equal(got["sums"][:, 0], f["sums"][:, 0])
    assert np.array_equal(got["sums"][ok], f["sums"][ok])
    assert np.array_equal(got["scaled"], f["scaled"])
    assert np.array_equal(got["deltas"], f["deltas"])
    assert np.array_equal(got["rates"], f["rates"], equal_nan=True)
    assert np.array_equal(got["ratios"], f["ratio_out"], equal_nan=True)


@pytest.mark.gpu
@pytest.mark.parametrize("case", DECODE_CASES)
def test_cuda_matches_reference_vectors(built, case):
    import torch
    f = np.load(os.path.join(GOLDEN, case + ".npz"))
    d = load_desc(f)
    tail, head = (int(x) for x in f["window"])
    ring = f["ring"]
    want = want_of(f)
    ctx = hbt.Context(0)
    try:
        ctx.ring_configure(1, 0)
        ctx.set_cpu_base(int(f["cpu"]))
        # seed the generator state the vector was produced with
        ctx.reset_state(int(f["carry_in_tstamp"]))
        if "carry_in_last" in f:
            ctx.set_carry(0, int(f["carry_in_tstamp"]), f["carry_in_last"])
        if d.aggregate == hbt.AGG_IP_HISTOGRAM:
            ctx.hist_configure(1 << 12)
        img = torch.from_numpy(ring.view(np.int64)).cuda()
        ctx.submit_ring_dev(0, img, ring.nbytes, tail, head)
        ctx.decode(d)
        got = ctx.fetch_results()[0]
        assert int(got["ret"]) == want["ret"]
        assert int(got["n_records"]) == want["n_records"]
        assert int(got["new_tail"]) == want["new_tail"]
        assert [int(x) for x in got["counters"]] == want["counters"]
        if d.aggregate == hbt.AGG_COUNT_DELTAS:
            assert int(got["n_rows"]) == want["n_rows"]
            assert int(got["num_unknown_count_interval"]) == want["num_unknown"]
            assert int(got["carry_tstamp"]) == want["carry_tstamp"]
            assert np.array_equal(ctx.fetch_rows(0, want["n_rows"], d.n_events), want["rows"])
            assert np.array_equal(ctx.fetch_carry(0, d.n_events), want["carry_last"])
        if "hist" in want:
            assert orc.hist_to_dict(ctx.hist_fetch()) == want["hist"]
    finally:
        ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("case", COUNTS_CASES)
def test_cuda_counts_match_reference_vectors(built, case):
    f = np.load(os.path.join(GOLDEN, case + ".npz"))
    ratios = [tuple(int(x) for x in r) for r in f["ratios"]]
    ctx = hbt.Context(0)
    try:
        got = ctx.counts_pipeline_host(f["reads"], ratios)
    finally:
        ctx.close()
    ok = f["sums"][:, 0] != 0
    assert np.array_equal(got["sums"][:, 0], f["sums"][:, 0])
    assert np.array_equal(got["sums"][ok], f["sums"][ok])
    assert np.array_equal(got["scaled"], f["scaled"])
    assert np.array_equal(got["deltas"], f["deltas"])
    finite = np.isfinite(f["rates"])
    assert np.array_equal(got["rates"][finite], f["rates"][finite])
    assert np.array_equal(np.isnan(got["rates"]), np.isnan(f["rates"]))
    fin = np.isfinite(f["ratio_out"])
    # north star: f64 ratios within 1e-12 relative
    assert np.allclose(got["ratios"][fin], f["ratio_out"][fin], rtol=1e-12, atol=0.0)


def check_switch(f, results, events):
    """results/events: per CPU result dicts and tid-tagged events (oracle port or CUDA path);
    mapped tid -> vid they must equal what the reference's own generator wrote."""
    mapped = orc.map_tids_to_vids(events)
    off = np.concatenate([[0], np.cumsum(f["n_events"])]).astype(int)
    for c in range(len(events)):
        r = results[c]
        assert r["ret"] == int(f["ret"][c]) and r["n_records"] == int(f["n_records"][c]), c
        assert r["new_tail"] == int(f["new_tail"][c]), c
        assert r["counters"] == [int(x) for x in f["counters"][c]], c
        assert r["carry_tstamp"] == int(f["last_tstamp"][c]), c
        assert r["num_unknown"] == int(f["in_write_errors"][c]), c
        want = f["events"][off[c]:off[c + 1]]
        assert len(mapped[c]) == len(want) and (mapped[c] == want).all(), c


@pytest.mark.parametrize("case", SWITCH_CASES)
def test_oracle_switch_events_match_reference_vectors(built, case):
    f = np.load(os.path.join(GOLDEN, case + ".npz"))
    d = hbt.switch_desc(tid_level=int(f["tid_level"]))
    res = [orc.port_switch_events(f["rings"][c], int(f["tails"][c]), int(f["heads"][c]), d,
                                  cpu=int(f["cpu_base"]) + c) for c in range(len(f["tails"]))]
    check_switch(f, res, [r["events"] for r in res])


@pytest.mark.gpu
@pytest.mark.parametrize("case", SWITCH_CASES)
def test_cuda_switch_events_match_reference_vectors(built, case):
    f = np.load(os.path.join(GOLDEN, case + ".npz"))
    n = len(f["tails"])
    ctx = hbt.Context(0)
    ctx.ring_configure(n, f["rings"][0].nbytes)
    ctx.set_cpu_base(int(f["cpu_base"]))
    for c in range(n):
        ctx.submit_ring_image(c, f["rings"][c], int(f["tails"][c]), int(f["heads"][c]))
    ctx.decode(hbt.switch_desc(tid_level=int(f["tid_level"])))
    res = ctx.fetch_results()
    results = [dict(ret=int(r["ret"]), n_records=int(r["n_records"]), new_tail=int(r["new_tail"]),
                    counters=[int(x) for x in r["counters"]], carry_tstamp=int(r["carry_tstamp"]),
                    num_unknown=int(r["num_unknown_count_interval"])) for r in res]
    events = [ctx.fetch_switch_events(c, int(res[c]["n_rows"])) for c in range(n)]
    ctx.close()
    check_switch(f, results, events)
