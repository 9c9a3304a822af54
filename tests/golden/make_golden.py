#!/usr/bin/env python
"""Generate tests/golden/*.npz: small input/output vectors produced by the
REFERENCE's own code (oracle/_ref/libhbt_ref.so, i.e. the reference headers under
/root/reference driven by oracle/ref_harness.cpp).

Run in the build container (where /root/reference exists):

    python tests/golden/make_golden.py

The committed .npz files let tests/test_golden.py pin the C restatement
(oracle/hbt_oracle.c) and the CUDA path to the reference on machines that have
neither the reference tree nor oracle/_ref.  Inputs are stored too, so the
vectors do not depend on the trace generator staying unchanged.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import dynolog_1_b200 as hbt  # noqa: E402
from dynolog_1_b200 import tracegen as tg  # noqa: E402
import _oracle as orc  # noqa: E402


def ring_np(stream, ds, tail, fill=77):
    ring, head = tg.place_in_ring(stream, ds, tail, fill_seed=fill)
    return ring.numpy().view(np.uint64), head


def desc_fields(d):
    return np.array([d.handled_mask, d.sample_layout, d.aggregate, d.n_events, d.ip_sample_bytes,
                     d.max_records, d.tsc_transform_mask, d.tsc.time_zero, d.tsc.time_mult,
                     d.tsc.time_shift], dtype=np.uint64)


def save_decode(name, ring, tail, head, d, res, carry_in=(hbt.INVALID_TS, None), cpu=0):
    out = dict(ring=ring, window=np.array([tail, head], dtype=np.uint64), desc=desc_fields(d),
               ret=np.int64(res["ret"]), n_records=np.uint64(res["n_records"]),
               new_tail=np.uint64(res["new_tail"]), n_rows=np.uint64(res["n_rows"]),
               counters=np.array(res["counters"], dtype=np.uint64),
               num_unknown=np.uint64(res["num_unknown"]),
               carry_tstamp=np.uint64(res["carry_tstamp"]),
               carry_in_tstamp=np.uint64(carry_in[0]), cpu=np.uint32(cpu))
    if carry_in[1] is not None:
        out["carry_in_last"] = np.asarray(carry_in[1], dtype=np.uint64)
    if "rows" in res:
        out["rows"] = res["rows"]
        out["carry_last"] = res["carry_last"]
    if res.get("hist") is not None:
        keys = sorted(res["hist"])
        out["hist_keys"] = np.array([[k[0], k[1], k[2]] for k in keys], dtype=np.uint64).reshape(-1, 3)
        out["hist_counts"] = np.array([res["hist"][k] for k in keys], dtype=np.uint64)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(f"{name}: ret={res['ret']} n_records={res['n_records']} n_rows={res['n_rows']}")


def main():
    assert orc.have_ref(), "oracle/_ref/libhbt_ref.so missing: run `make -C oracle ref` first"
    ds = 1 << 15

    # --- CpuCountSampleGenerator deltas -------------------------------------------
    E = 4
    s = tg.group_sample_stream(101, 400, E, lost_every=37)
    tail = 3 * ds + ds - 4096 - 8 * 5       # the window wraps the ring end
    ring, head = ring_np(s, ds, tail)
    d = hbt.sampling_desc(E)
    save_decode("deltas_lost", ring, tail, head, d, orc.ref_consume(ring, tail, head, d))
    d = hbt.sampling_desc(E, max_records=123)
    carry = (999_000, np.array([5, 6, 7, 8], dtype=np.uint64))
    save_decode("deltas_max_records_carry", ring, tail, head, d,
                orc.ref_consume(ring, tail, head, d, carry[0], carry[1]), carry)
    s = tg.group_sample_stream(102, 300, E, regress_at=211)
    ring, head = ring_np(s, ds, 8 * 77)
    d = hbt.sampling_desc(E)
    save_decode("deltas_tstamp_regression", ring, 8 * 77, head, d,
                orc.ref_consume(ring, 8 * 77, head, d))
    s = tg.group_sample_stream(103, 300, E)
    s[(5 + E) * 150 + 2] = E + 1
    ring, head = ring_np(s, ds, 0)
    save_decode("deltas_bad_sample_size", ring, 0, head, d, orc.ref_consume(ring, 0, head, d))
    s = tg.group_sample_stream(104, 350, 6, lost_every=90)
    ring, head = ring_np(s, ds, 4096 + 64)
    d = hbt.sampling_desc(6, tsc_mask=0b100101, tsc=(987654321, 1234567, 21))
    save_decode("deltas_tsc_transform", ring, 4096 + 64, head, d,
                orc.ref_consume(ring, 4096 + 64, head, d))
    s = tg.group_sample_stream(105, 60, 32, lost_every=11)
    ring, head = ring_np(s, ds, 5 * ds + 3000 * 8)
    d = hbt.sampling_desc(32)
    save_decode("deltas_32_events", ring, 5 * ds + 3000 * 8, head, d,
                orc.ref_consume(ring, 5 * ds + 3000 * 8, head, d))

    # --- consume() counters per mode over a mixed stream --------------------------------
    # every record type consume() dispatches on, THROTTLE / UNTHROTTLE / READ / AUX / ITRACE_START
    # included (CpuEventsGroup.h:1409-1420, 1446-1496)
    s = tg.mixed_stream(201, 330, 8, mix=tg.FULL_MIX)
    tail = 7 * ds + ds - 2048
    ring, head = ring_np(s, ds, tail)
    for name, mode in (("sampling", hbt.MODE_SAMPLING), ("ctxswitch", hbt.MODE_CONTEXT_SWITCH),
                       ("auxspace", hbt.MODE_AUX_SPACE), ("mixed", hbt.MODE_MIXED)):
        d = hbt.count_desc(mode)
        save_decode(f"count_{name}", ring, tail, head, d, orc.ref_consume(ring, tail, head, d))
    d = hbt.count_desc(hbt.MODE_CONTEXT_SWITCH, max_records=100)
    save_decode("count_ctxswitch_max100", ring, tail, head, d, orc.ref_consume(ring, tail, head, d))
    # a READ whose size disagrees with its nr: -EPERM in front of it (CpuEventsGroup.h:1446-1456)
    w = s.clone()
    pos, k = 0, 0
    while pos < w.numel():
        h = int(w[pos].item()) & ((1 << 64) - 1)
        if (h & 0xFFFFFFFF) == tg.REC_READ:
            k += 1
            if k == 5:
                w[pos + 2] = 9
                break
        pos += (h >> 48) // 8
    assert k == 5
    ring, head = ring_np(w, ds, tail)
    for name, mode in (("sampling", hbt.MODE_SAMPLING), ("ctxswitch", hbt.MODE_CONTEXT_SWITCH)):
        d = hbt.count_desc(mode)
        save_decode(f"count_{name}_bad_read", ring, tail, head, d, orc.ref_consume(ring, tail, head, d))
    ring, head = ring_np(s, ds, tail)
    mix = (("sample", 0.6), ("switch", 0.2), ("comm", 0.05), ("mmap", 0.05), ("lost", 0.03),
           ("fork", 0.035), ("exit", 0.035), ("read", 0.03), ("throttle", 0.01), ("unthrottle", 0.01),
           ("aux", 0.01))
    s = tg.mixed_stream(202, 330, 8, mix=mix)
    ring, head = ring_np(s, ds, tail)
    d = hbt.sampling_desc(8, handled=hbt.MODE_MIXED)
    save_decode("mixed_deltas", ring, tail, head, d, orc.ref_consume(ring, tail, head, d))

    # --- IP samples: consume() + the harness's (pid,ip) map (no reference handler) -------
    s = tg.ip_sample_stream(301, 700, cpu=5, n_pids=16, n_ips=64, lost_every=97)
    tail = 2 * ds + ds - 1024
    ring, head = ring_np(s, ds, tail)
    d = hbt.ip_desc(40)
    save_decode("ip_hist_40", ring, tail, head, d, orc.ref_consume(ring, tail, head, d, cpu=5), cpu=5)

    # --- counter groups: PerCpuBase::read + getCount + PerfMonitor::log math --------------
    reads = tg.counts_trace(401, 24, 8, 4, multiplexed=True).numpy().view(np.uint64).copy()
    reads[3, :, 3] = np.uint64(2**62)
    reads[5, :, 4] = np.uint64(2**60 + 12345)
    reads[7, :, 1] = 0
    reads[7, :, 2] = 0
    reads[9, 2, 0] = 0
    ratios = [(1, 0), (2, 1), (3, 0)]
    o = orc.ref_counts(reads, ratios)
    np.savez_compressed(os.path.join(HERE, "counts_mux.npz"), reads=reads,
                        ratios=np.array(ratios, dtype=np.uint32), sums=o["sums"], scaled=o["scaled"],
                        rates=o["rates"], ratio_out=o["ratios"], deltas=o["deltas"])
    reads = tg.counts_trace(402, 16, 64, 8, multiplexed=False).numpy().view(np.uint64)
    o = orc.ref_counts(reads, [(1, 0)])
    np.savez_compressed(os.path.join(HERE, "counts_c1_shape.npz"), reads=reads,
                        ratios=np.array([(1, 0)], dtype=np.uint32), sums=o["sums"],
                        scaled=o["scaled"], rates=o["rates"], ratio_out=o["ratios"],
                        deltas=o["deltas"])
    print("counts: ok")

    # --- context switches: the real CpuThreadSwitchGenerator (vid-tagged tagstack::Events) ---
    for name, seed, lost_every, n_rec in (("switch_events_lost", 5, 23, 500), ("switch_events_plain", 6, 0, 300)):
        dsz = 1 << 15
        rings, tails, heads = [], [], []
        for c in range(3):
            st = tg.ctx_switch_stream_consistent(seed * 100 + c, n_rec, lost_every=lost_every if c != 1 else 0,
                                                 first_tid=1000 + 10000 * c)
            tail = 3 * dsz + dsz - 1024 - 8 * c
            ring, head = ring_np(st, dsz, tail)
            rings.append(ring)
            tails.append(tail)
            heads.append(head)
        rings = np.stack(rings)
        ref = orc.ref_switch_events(rings, tails, heads, tid_level=2, cpu_base=8)
        np.savez_compressed(
            os.path.join(HERE, name + ".npz"), rings=rings, tails=np.array(tails, np.uint64),
            heads=np.array(heads, np.uint64), tid_level=np.uint32(2), cpu_base=np.uint32(8),
            events=np.concatenate([r["events"] for r in ref]),
            n_events=np.array([len(r["events"]) for r in ref], np.uint64),
            ret=np.array([r["ret"] for r in ref], np.int64),
            n_records=np.array([r["n_records"] for r in ref], np.uint64),
            new_tail=np.array([r["new_tail"] for r in ref], np.uint64),
            counters=np.array([r["counters"] for r in ref], np.uint64),
            last_tstamp=np.array([r["carry_tstamp"] for r in ref], np.uint64),
            in_write_errors=np.array([r["num_unknown"] for r in ref], np.uint64))
        print(f"{name}: events per cpu {[len(r['events']) for r in ref]}")


if __name__ == "__main__":
    main()
