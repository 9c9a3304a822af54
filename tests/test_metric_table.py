"""The library's static metric-descriptor table against the reference's own registry:
tests/golden/builtin_metrics.json is what makeAvailableMetrics() (BuiltinMetrics.cpp:470-1175)
holds, dumped by oracle/dump_builtin_metrics.cpp compiled against the reference sources
(tests/golden/make_builtin_metrics.py)."""
import json
import os

import numpy as np
import pytest

import dynolog_1_b200 as hbt
from dynolog_1_b200 import tracegen as tg
import _oracle as orc

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "builtin_metrics.json")


def test_table_equals_reference_registry(built):
    ref = json.load(open(GOLDEN))
    table = hbt.metric_table()
    builtin = [d for d in table if d["builtin"]]
    # every (metric id, architecture) of the registry, nothing else, events in the same order
    want = {(mid, arch): [e["nickname"] for e in evs] for mid, by_arch in ref.items()
            for arch, evs in by_arch.items()}
    got = {(d["id"], d["arch"]): d["nicknames"] for d in builtin}
    assert got == want
    assert len(ref) == 19
    # BuiltinMetricsTest.cpp:34-41: ipc is made of exactly two (generic) events
    assert got[("ipc", "any")] == ["cycles", "instructions"]
    # value transforms: topdown_l1's unhalted_time is the one TscToKernelTime event (PmuEvent.h:218-226)
    for d in builtin:
        m = 0
        for i, e in enumerate(ref[d["id"]][d["arch"]]):
            m |= e["tsc_to_kernel_time"] << i
        assert d["tsc_transform_mask"] == m, d["id"]
    assert [d["index"] for d in table] == list(range(len(table)))


def test_named_ratios_refer_to_the_events_they_name(built):
    table = {(d["id"], d["arch"]): d for d in hbt.metric_table()}
    ipc = table[("ipc", "any")]
    assert ipc["ratios"] == [("ipc", ipc["nicknames"].index("instructions"), ipc["nicknames"].index("cycles"))]
    l3 = table[("l3_cache_misses_per_instruction", "any")]
    assert l3["ratios"] == [("l3_cache_misses_per_instruction", 0, 1)]
    for name in ("llc_load_miss_rate", "dtlb_load_miss_rate"):   # build-defined groups
        d = table[(name, "any")]
        assert not d["builtin"] and d["ratios"] == [(name, 0, 1)]
    for d in table.values():
        for _, a, b in d["ratios"]:
            assert a < len(d["nicknames"]) and b < len(d["nicknames"])


def test_find_follows_get_event_refs(built):
    """MetricDesc::getEventRefs (Metrics.h:162-180): the architecture's own list first, then the
    one registered for any architecture, else nothing."""
    lib = hbt.load_library()
    d = hbt.metric_find("topdown_l1", "SKX")
    assert d and d.contents.arch == b"SKX" and d.contents.n_events == 6
    assert hbt.metric_find("topdown_l1", "ICL") is None and hbt.metric_find("topdown_l1") is None
    d = hbt.metric_find("ipc", "SKX")
    assert d and d.contents.arch == b"any"
    assert hbt.metric_find("no_such_metric") is None
    assert lib.hbt_metric_event_index(d, b"instructions") == 1   # eventIdxFromNicknames
    assert lib.hbt_metric_event_index(d, b"nope") < 0
    assert not lib.hbt_metric_at(lib.hbt_metric_count())


@pytest.mark.gpu
@pytest.mark.parametrize("metric", ["ipc", "cs_ipc", "l3_cache_misses_per_instruction", "dqos",
                                    "llc_load_miss_rate", "instructions"])
def test_pipeline_selected_by_metric_id(built, metric):
    import torch
    d = hbt.metric_find(metric)
    E, R = d.contents.n_events, d.contents.n_ratios
    I, Cc = 300, 48
    reads = tg.counts_trace(900 + E, I, Cc, E, device="cuda", multiplexed=True)
    ratios = [(d.contents.ratios[k].num_event, d.contents.ratios[k].den_event) for k in range(R)]
    want = orc.counts(reads.cpu().numpy().view(np.uint64), ratios)
    sums = torch.zeros((I, 3 + E), dtype=torch.int64, device="cuda")
    scaled = torch.zeros((I, E), dtype=torch.int64, device="cuda")
    rates = torch.zeros((I, E), dtype=torch.float32, device="cuda")
    ratio = torch.zeros((I, max(R, 1)), dtype=torch.float64, device="cuda")
    deltas = torch.zeros((I, E), dtype=torch.int64, device="cuda")
    ctx = hbt.Context(0)
    try:
        ctx.counts_pipeline_metric_dev(d, reads, I, Cc, sums, scaled, rates, ratio if R else None, deltas)
        ctx.sync()
    finally:
        ctx.close()
    assert np.array_equal(sums.cpu().numpy().view(np.uint64), want["sums"])
    assert np.array_equal(scaled.cpu().numpy().view(np.uint64), want["scaled"])
    assert np.array_equal(deltas.cpu().numpy().view(np.uint64), want["deltas"])
    assert np.array_equal(rates.cpu().numpy(), want["rates"])
    if R:   # f64 ratios: within 1e-12 relative (north star); in fact identical (IEEE div)
        assert np.allclose(ratio.cpu().numpy()[:, :R], want["ratios"], rtol=1e-12, atol=0.0)
