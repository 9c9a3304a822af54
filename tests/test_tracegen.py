"""The synthetic trace generator: deterministic, device independent, well formed."""
import numpy as np
import torch

import dynolog_1_b200 as hbt
from dynolog_1_b200 import tracegen as tg
import _oracle as orc


def test_splitmix64_known_answers():
    # splitmix64 reference outputs for state 0 (Vigna): next() adds the gamma first
    x = torch.tensor([0], dtype=torch.int64)
    v = int(tg.splitmix64(x).item()) & (2**64 - 1)
    assert v == 0xE220A8397B1DCDAF
    v2 = int(tg.splitmix64(torch.tensor([tg._s64(0x9E3779B97F4A7C15)])).item()) & (2**64 - 1)
    assert v2 == 0x6E789E6AA1B965F4


def test_streams_are_deterministic_and_seed_dependent():
    a = tg.group_sample_stream(7, 100, 8, lost_every=13)
    b = tg.group_sample_stream(7, 100, 8, lost_every=13)
    c = tg.group_sample_stream(8, 100, 8, lost_every=13)
    assert torch.equal(a, b) and not torch.equal(a, c)
    assert tg.stream_seed(2, 5) == 0x5EED000000000000 + (2 << 32) + 5


def test_place_in_ring_wraps():
    s = torch.arange(1, 101, dtype=torch.int64)
    ring, head = tg.place_in_ring(s, 1024, tail=1024 * 3 + 1000)
    r = ring.numpy()
    assert head == 1024 * 3 + 1000 + 800
    assert r[125] == 1 and r[127] == 3 and r[0] == 4 and r[96] == 100


def test_mixed_stream_is_well_formed(built):
    s = tg.mixed_stream(3, 3000, 8)
    ring, head = tg.place_in_ring(s, 1 << 19, 0)
    d = hbt.count_desc(hbt.MODE_MIXED)
    r = orc.port_consume(ring.numpy().view(np.uint64), 0, head, d)
    assert r["ret"] == -61 and r["n_records"] == 3000 and r["new_tail"] == head
    c = r["counters"]
    assert c[hbt.CNT_SAMPLE] > 1800 and c[hbt.CNT_SWITCH_CPU_WIDE] > 300 and c[hbt.CNT_COMM] > 80


def test_counts_trace_invariants():
    t = tg.counts_trace(1, 50, 16, 8, multiplexed=True).numpy()
    assert (t[:, :, 0] == 8).all()
    assert (t[:, :, 2] <= t[:, :, 1]).all() and (t[:, :, 2] > 0).all()
    assert (np.diff(t[:, :, 1:], axis=0) > 0).all()  # cumulative since enable()
