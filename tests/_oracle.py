"""ctypes access to the CHECKERS under oracle/ (test infrastructure only):
liboracle.so (plain-C restatement) and _ref/libhbt_ref.so (the reference's own
headers driven by oracle/ref_harness.cpp)."""
import ctypes as C
import os

import numpy as np

import dynolog_1_b200 as hbt

ORACLE_DIR = os.path.join(hbt.REPO_ROOT, "oracle")
PORT_PATH = os.path.join(ORACLE_DIR, "liboracle.so")
REF_PATH = os.path.join(ORACLE_DIR, "_ref", "libhbt_ref.so")

_P = C.c_void_p


class Carry(C.Structure):
    _fields_ = [("tstamp", C.c_uint64), ("last", C.c_uint64 * hbt.MAX_EVENTS)]


def _np(a):
    return C.c_void_p(a.ctypes.data) if a is not None else None


_port = None
_ref = None


def port():
    global _port
    if _port is None:
        lib = C.CDLL(PORT_PATH)
        lib.hbt_oracle_consume.restype = C.c_int
        lib.hbt_oracle_consume.argtypes = [_P, C.c_uint64, C.c_uint64, C.c_uint64,
                                           C.POINTER(hbt.DecodeDesc), C.POINTER(Carry), _P,
                                           C.c_uint64, _P, C.POINTER(hbt.CpuResult)]
        lib.hbt_oracle_counts.restype = C.c_int
        lib.hbt_oracle_counts.argtypes = [_P, C.c_uint64, C.c_uint32, C.c_uint32, _P, C.c_uint32,
                                          _P, _P, _P, _P, _P]
        lib.hbt_oracle_hist_new.restype = _P
        lib.hbt_oracle_hist_free.argtypes = [_P]
        lib.hbt_oracle_hist_size.restype = C.c_uint64
        lib.hbt_oracle_hist_size.argtypes = [_P]
        lib.hbt_oracle_hist_fetch.restype = C.c_uint64
        lib.hbt_oracle_hist_fetch.argtypes = [_P, C.c_uint32, _P, C.c_uint64]
        lib.hbt_oracle_kernel_time_from_tsc.restype = C.c_uint64
        lib.hbt_oracle_kernel_time_from_tsc.argtypes = [C.c_uint64, C.POINTER(hbt.TscParams)]
        _port = lib
    return _port


def have_ref():
    return os.path.exists(REF_PATH)


def ref():
    global _ref
    if _ref is None:
        lib = C.CDLL(REF_PATH)
        lib.hbtref_consume_count.restype = C.c_int
        lib.hbtref_consume_count.argtypes = [_P, C.c_uint64, C.c_uint64, C.c_uint64, C.c_int,
                                             C.c_uint32, C.POINTER(hbt.CpuResult)]
        for name in ("hbtref_sample_deltas", "hbtref_mixed_deltas"):
            fn = getattr(lib, name)
            fn.restype = C.c_int
            fn.argtypes = [_P, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint64,
                           C.POINTER(hbt.TscParams), C.c_uint32, C.POINTER(C.c_uint64), _P, _P,
                           C.c_uint64, C.POINTER(hbt.CpuResult)]
        lib.hbtref_hist_new.restype = _P
        lib.hbtref_hist_free.argtypes = [_P]
        lib.hbtref_ip_hist.restype = C.c_double
        lib.hbtref_ip_hist.argtypes = [_P, _P, C.c_uint64, _P, _P, C.c_uint32, C.c_uint32,
                                       C.c_uint32, C.c_uint32, _P]
        lib.hbtref_hist_size.restype = C.c_uint64
        lib.hbtref_hist_size.argtypes = [_P]
        lib.hbtref_hist_fetch.restype = C.c_uint64
        lib.hbtref_hist_fetch.argtypes = [_P, _P, C.c_uint64]
        lib.hbtref_sample_deltas_timed.restype = C.c_double
        lib.hbtref_sample_deltas_timed.argtypes = [_P, C.c_uint64, _P, _P, C.c_uint32, C.c_uint32,
                                                   C.c_uint32, C.POINTER(C.c_uint64),
                                                   C.POINTER(C.c_uint64)]
        lib.hbtref_counts.restype = C.c_double
        lib.hbtref_counts.argtypes = [_P, C.c_uint64, C.c_uint32, C.c_uint32, _P, C.c_uint32, _P,
                                      _P, _P, _P, _P, C.c_uint32]
        lib.hbtref_kernel_time_from_tsc.restype = C.c_uint64
        lib.hbtref_kernel_time_from_tsc.argtypes = [C.c_uint64, C.POINTER(hbt.TscParams)]
        _ref = lib
    return _ref


MODE_INDEX = {hbt.MODE_SAMPLING: 0, hbt.MODE_CONTEXT_SWITCH: 1, hbt.MODE_AUX_SPACE: 2,
              hbt.MODE_MIXED: 3}


def result_dict(res: hbt.CpuResult, rows=None, carry_last=None, hist=None):
    d = dict(ret=int(res.ret), n_records=int(res.n_records), new_tail=int(res.new_tail),
             n_rows=int(res.n_rows), counters=[int(x) for x in res.counters],
             num_unknown=int(res.num_unknown_count_interval),
             carry_tstamp=int(res.carry_tstamp))
    if rows is not None:
        d["rows"] = rows
    if carry_last is not None:
        d["carry_last"] = carry_last
    if hist is not None:
        d["hist"] = hist
    return d


def hist_to_dict(entries: np.ndarray):
    return {(int(e["cpu"]), int(e["pid"]), int(e["ip"])): int(e["count"]) for e in entries}


def port_consume(ring: np.ndarray, tail: int, head: int, desc: hbt.DecodeDesc,
                 carry_tstamp=hbt.INVALID_TS, carry_last=None, counters=None, num_unknown=0,
                 cpu=0, rows_budget=None):
    """oracle/hbt_oracle.c over one ring image (uint8/uint64 numpy array).  rows_budget: rows the
    generator's output ring still takes (None: unlimited)."""
    lib = port()
    ring = np.ascontiguousarray(ring).view(np.uint8)
    n = max(1, desc.n_events)
    carry = Carry()
    carry.tstamp = carry_tstamp
    if carry_last is not None:
        for i, v in enumerate(carry_last):
            carry.last[i] = int(v)
    res = hbt.CpuResult()
    if counters is not None:
        for i, v in enumerate(counters):
            res.counters[i] = int(v)
    res.num_unknown_count_interval = num_unknown
    cap = (head - tail) // (40 + 8 * n) + 2
    rows = np.zeros((cap, 2 + n), dtype=np.uint64)
    if rows_budget is not None:
        cap = min(cap, rows_budget)
    h = lib.hbt_oracle_hist_new() if desc.aggregate == hbt.AGG_IP_HISTOGRAM else None
    rc = lib.hbt_oracle_consume(_np(ring), ring.nbytes, tail, head, C.byref(desc), C.byref(carry),
                                _np(rows), cap, h, C.byref(res))
    assert rc == 0
    hist = None
    if h:
        m = lib.hbt_oracle_hist_size(h)
        ent = np.zeros(max(1, m), dtype=hbt.HIST_DTYPE)
        k = lib.hbt_oracle_hist_fetch(h, cpu, _np(ent), ent.shape[0])
        hist = hist_to_dict(ent[:k])
        lib.hbt_oracle_hist_free(h)
    return result_dict(res, rows[:res.n_rows].copy(),
                       np.array([carry.last[i] for i in range(n)], dtype=np.uint64), hist)


def ref_consume(ring: np.ndarray, tail: int, head: int, desc: hbt.DecodeDesc,
                carry_tstamp=hbt.INVALID_TS, carry_last=None, counters=None, num_unknown=0,
                cpu=0):
    """The reference's own consume() (oracle/_ref) over one ring image."""
    lib = ref()
    ring = np.ascontiguousarray(ring).view(np.uint8)
    res = hbt.CpuResult()
    if counters is not None:
        for i, v in enumerate(counters):
            res.counters[i] = int(v)
    res.num_unknown_count_interval = num_unknown
    if desc.aggregate == hbt.AGG_NONE and desc.sample_layout == hbt.SAMPLE_GROUP:
        rc = lib.hbtref_consume_count(_np(ring), ring.nbytes, tail, head,
                                      MODE_INDEX[desc.handled_mask], desc.max_records,
                                      C.byref(res))
        assert rc == 0
        res.carry_tstamp = hbt.INVALID_TS
        return result_dict(res)
    if desc.aggregate == hbt.AGG_COUNT_DELTAS:
        n = desc.n_events
        ts = C.c_uint64(carry_tstamp)
        last = np.zeros(hbt.MAX_EVENTS, dtype=np.uint64)
        if carry_last is not None:
            last[:len(carry_last)] = carry_last
        cap = (head - tail) // (40 + 8 * n) + 2
        rows = np.zeros((cap, 2 + n), dtype=np.uint64)
        fn = lib.hbtref_sample_deltas if desc.handled_mask == hbt.MODE_SAMPLING \
            else lib.hbtref_mixed_deltas
        rc = fn(_np(ring), ring.nbytes, tail, head, n, desc.tsc_transform_mask,
                C.byref(desc.tsc), desc.max_records, C.byref(ts), _np(last), _np(rows), cap,
                C.byref(res))
        assert rc == 0, rc
        return result_dict(res, rows[:res.n_rows].copy(), last[:n].copy())
    # IP layouts
    h = lib.hbtref_hist_new()
    tails = np.array([tail], dtype=np.uint64)
    heads = np.array([head], dtype=np.uint64)
    resarr = (hbt.CpuResult * 1)()
    if counters is not None:
        for i, v in enumerate(counters):
            resarr[0].counters[i] = int(v)
    t = lib.hbtref_ip_hist(h, _np(ring), ring.nbytes, _np(tails), _np(heads), 1,
                           desc.ip_sample_bytes, desc.max_records, 1, resarr)
    assert t >= 0, t
    m = lib.hbtref_hist_size(h)
    ent = np.zeros(max(1, m), dtype=hbt.HIST_DTYPE)
    k = lib.hbtref_hist_fetch(h, _np(ent), ent.shape[0])
    ent = ent[:k]
    ent["cpu"] = cpu
    hist = hist_to_dict(ent) if desc.aggregate == hbt.AGG_IP_HISTOGRAM else None
    lib.hbtref_hist_free(h)
    resarr[0].carry_tstamp = hbt.INVALID_TS
    return result_dict(resarr[0], hist=hist)


def ref_consume_budget(ring, tail, head, desc, rows_budget, carry_tstamp=hbt.INVALID_TS, carry_last=None,
                       counters=None, num_unknown=0):
    """The reference's own CpuCountSampleGenerator with an output ring that has room for exactly
    rows_budget more rows (oracle/_ref, hbtref_sample_deltas_budget)."""
    lib = ref()
    ring = np.ascontiguousarray(ring).view(np.uint8)
    n = desc.n_events
    res = hbt.CpuResult()
    if counters is not None:
        for i, v in enumerate(counters):
            res.counters[i] = int(v)
    res.num_unknown_count_interval = num_unknown
    ts = C.c_uint64(carry_tstamp)
    last = np.zeros(hbt.MAX_EVENTS, dtype=np.uint64)
    if carry_last is not None:
        last[:len(carry_last)] = carry_last
    cap = (head - tail) // (40 + 8 * n) + 2
    rows = np.zeros((cap, 2 + n), dtype=np.uint64)
    lib.hbtref_sample_deltas_budget.restype = C.c_int
    lib.hbtref_sample_deltas_budget.argtypes = [_P, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32,
                                                C.c_uint64, C.POINTER(C.c_uint64), _P, _P, C.c_uint64,
                                                C.POINTER(hbt.CpuResult)]
    with quiet_stderr():
        rc = lib.hbtref_sample_deltas_budget(_np(ring), ring.nbytes, tail, head, n, desc.max_records, rows_budget,
                                             C.byref(ts), _np(last), _np(rows), cap, C.byref(res))
    assert rc == 0, rc
    return result_dict(res, rows[:res.n_rows].copy(), last[:n].copy())


def _ref_covers(desc: hbt.DecodeDesc) -> bool:
    """Decodes ref_consume can run through the reference's own classes."""
    if desc.aggregate == hbt.AGG_NONE and desc.sample_layout == hbt.SAMPLE_GROUP:
        return desc.handled_mask in MODE_INDEX
    if desc.aggregate == hbt.AGG_COUNT_DELTAS:
        return desc.handled_mask in (hbt.MODE_SAMPLING, hbt.MODE_MIXED)
    if desc.sample_layout == hbt.SAMPLE_IP and desc.aggregate in (hbt.AGG_NONE, hbt.AGG_IP_HISTOGRAM):
        return desc.handled_mask == (hbt.H_LOST | hbt.H_SAMPLE) and desc.ip_sample_bytes in IP_REF_SIZES
    return False


# hbtref_ip_hist instantiates its IP sample mode for these record sizes only
IP_REF_SIZES = (40, 48, 64, 104, 128)


def consume(ring: np.ndarray, tail: int, head: int, desc: hbt.DecodeDesc, carry_tstamp=hbt.INVALID_TS,
            carry_last=None, counters=None, num_unknown=0, cpu=0, rows_budget=None):
    """What the GPU tests compare the CUDA path with: the REFERENCE's own consume() (oracle/_ref)
    whenever that library is present and covers the decode -- the C restatement is then only
    cross-checked against it here -- else the C restatement alone."""
    p = port_consume(ring, tail, head, desc, carry_tstamp, carry_last, counters, num_unknown, cpu, rows_budget)
    if not (have_ref() and _ref_covers(desc)):
        return p
    if rows_budget is not None:
        assert desc.aggregate == hbt.AGG_COUNT_DELTAS and desc.handled_mask == hbt.MODE_SAMPLING and not desc.tsc_transform_mask
        r = ref_consume_budget(ring, tail, head, desc, rows_budget, carry_tstamp, carry_last, counters, num_unknown)
    else:
        with quiet_stderr():
            r = ref_consume(ring, tail, head, desc, carry_tstamp, carry_last, counters, num_unknown, cpu)
    for k in ("ret", "n_records", "new_tail", "counters"):
        assert r[k] == p[k], ("oracle port != reference", k, r[k], p[k])
    if desc.aggregate == hbt.AGG_COUNT_DELTAS:
        for k in ("n_rows", "num_unknown", "carry_tstamp"):
            assert r[k] == p[k], ("oracle port != reference", k, r[k], p[k])
        assert np.array_equal(r["rows"], p["rows"]) and np.array_equal(r["carry_last"], p["carry_last"])
    else:
        r["carry_tstamp"] = p["carry_tstamp"]
    if desc.aggregate == hbt.AGG_IP_HISTOGRAM:
        assert r["hist"] == p["hist"]
    r["checked_against"] = "reference"
    return r


class quiet_stderr:
    """The reference logs a line per LOST / THROTTLE / unknown record (CpuEventsGroup.h:1409-1420,
    1498-1500; PerCpuCountSampleGenerator.h:111): keep that out of the test log."""

    def __enter__(self):
        import sys
        sys.stderr.flush()
        self.saved, self.null = os.dup(2), os.open(os.devnull, os.O_WRONLY)
        os.dup2(self.null, 2)

    def __exit__(self, *a):
        os.dup2(self.saved, 2)
        os.close(self.saved)
        os.close(self.null)


def switch_events(rings, tails, heads, desc: hbt.DecodeDesc, cpu_base=0, last_tstamp=None,
                  in_write_errors=None, counters=None):
    """Per-CPU oracle answer for an AGG_SWITCH_EVENTS decode (tid-tagged events of the C
    restatement).  With oracle/_ref present the REAL CpuThreadSwitchGenerator runs over the same
    rings (CPUs in ascending order, one shared thread map) and every field -- and the events after
    the tid -> vid mapping -- must equal its output; "ref_events" then holds its vid-tagged events
    for a direct comparison with the GPU's mapped stream.  Streams must be ones the reference's
    handlers accept (see tracegen.ctx_switch_stream)."""
    n = len(tails)
    out = []
    for c in range(n):
        out.append(port_switch_events(
            rings[c], int(tails[c]), int(heads[c]), desc, cpu=cpu_base + c,
            last_tstamp=0 if last_tstamp is None else int(last_tstamp[c]),
            in_write_errors=0 if in_write_errors is None else int(in_write_errors[c]),
            counters=None if counters is None else counters[c]))
    if have_ref() and desc.handled_mask == hbt.MODE_CONTEXT_SWITCH and counters is None:
        with quiet_stderr():
            ref_out = ref_switch_events(np.stack([np.asarray(r) for r in rings]), tails, heads,
                                        tid_level=desc.tid_level, cpu_base=cpu_base,
                                        max_records=desc.max_records, last_tstamp=last_tstamp,
                                        in_write_errors=in_write_errors)
        mapped = map_tids_to_vids([o["events"] for o in out])
        for c in range(n):
            for k in ("ret", "n_records", "new_tail", "counters", "carry_tstamp", "num_unknown"):
                assert ref_out[c][k] == out[c][k], ("oracle port != reference", c, k)
            assert len(ref_out[c]["events"]) == len(mapped[c]) and (ref_out[c]["events"] == mapped[c]).all()
            out[c]["ref_events"] = ref_out[c]["events"]
    return out


def port_switch_events(ring: np.ndarray, tail: int, head: int, desc: hbt.DecodeDesc, cpu=0,
                       last_tstamp=0, in_write_errors=0, counters=None):
    """oracle/hbt_oracle.c with CpuThreadSwitchGenerator's handlers over one ring image:
    tid-tagged events (SWITCH_EVENT_DTYPE, COMM pseudo-events included)."""
    lib = port()
    lib.hbt_oracle_consume_cpu.restype = C.c_int
    lib.hbt_oracle_consume_cpu.argtypes = [_P, C.c_uint64, C.c_uint64, C.c_uint64,
                                           C.POINTER(hbt.DecodeDesc), C.POINTER(Carry), _P, C.c_uint64,
                                           _P, C.POINTER(hbt.CpuResult), C.c_uint32]
    ring = np.ascontiguousarray(ring).view(np.uint8)
    carry = Carry()
    carry.tstamp = last_tstamp
    carry.last[0] = in_write_errors
    res = hbt.CpuResult()
    if counters is not None:
        for i, v in enumerate(counters):
            res.counters[i] = int(v)
    cap = 2 * ((head - tail) // 8 + 2)
    ev = np.zeros(cap, dtype=hbt.SWITCH_EVENT_DTYPE)
    rc = lib.hbt_oracle_consume_cpu(_np(ring), ring.nbytes, tail, head, C.byref(desc), C.byref(carry),
                                    _np(ev), cap, None, C.byref(res), cpu)
    assert rc == 0
    d = result_dict(res)
    d["events"] = ev[:res.n_rows].copy()
    return d


def ref_switch_events(rings: np.ndarray, tails, heads, tid_level=0, cpu_base=0, max_records=hbt.DRAIN,
                      last_tstamp=None, in_write_errors=None):
    """The reference's own CpuThreadSwitchGenerator over rings[n_cpus, data_size/8]: per CPU
    result dict with vid-tagged events, CPUs processed in ascending order sharing one
    PerThreadStats."""
    lib = ref()
    rings = np.ascontiguousarray(rings)
    n_cpus = rings.shape[0]
    ds = rings[0].nbytes
    tails = np.asarray(tails, dtype=np.uint64)
    heads = np.asarray(heads, dtype=np.uint64)
    lt = np.zeros(n_cpus, dtype=np.uint64) if last_tstamp is None else np.asarray(last_tstamp, np.uint64).copy()
    ie = np.zeros(n_cpus, dtype=np.uint8) if in_write_errors is None else np.asarray(in_write_errors, np.uint8).copy()
    cap = int(sum(2 * ((int(h) - int(t)) // 32 + 2) for t, h in zip(tails, heads)))
    ev = np.zeros(cap, dtype=hbt.SWITCH_EVENT_DTYPE)
    off = np.zeros(n_cpus + 1, dtype=np.uint64)
    res = (hbt.CpuResult * n_cpus)()
    lib.hbtref_switch_events.restype = C.c_int
    lib.hbtref_switch_events.argtypes = [_P, C.c_uint64, _P, _P, C.c_uint32, C.c_uint32, C.c_uint32,
                                         C.c_uint32, _P, _P, _P, C.c_uint64, _P, _P]
    rc = lib.hbtref_switch_events(_np(rings), ds, _np(tails), _np(heads), n_cpus, cpu_base, tid_level,
                                  max_records, _np(lt), _np(ie), _np(ev), cap, _np(off), res)
    assert rc == 0, rc
    out = []
    for c in range(n_cpus):
        d = result_dict(res[c])
        d["events"] = ev[int(off[c]):int(off[c + 1])].copy()
        out.append(d)
    return out


def map_tids_to_vids(per_cpu_events):
    """Test-side restatement of PerThreadStats' virtual-id bookkeeping as the handlers drive
    it (PerCpuThreadSwitchGenerator.h:52-135, 228-412): turns tid-tagged event streams
    (CPUs in ascending order, COMM pseudo-events included) into the reference's vid-tagged
    ones (pseudo-events dropped)."""
    next_vid = 1
    active = {}
    out = []
    for ev in per_cpu_events:
        keep = np.ones(len(ev), dtype=bool)
        ev = ev.copy()
        for i in range(len(ev)):
            t, tid = int(ev["type"][i]), int(ev["tag"][i])
            if t == hbt.SWEV_WRITE_ERRORS_START:
                active.clear()                                  # :253-256
                continue
            if t not in (hbt.SWEV_COMM, hbt.SWEV_THREAD_CREATION, hbt.SWEV_THREAD_DESTRUCTION,
                         hbt.SWEV_SWITCH_IN):
                continue
            if tid == 0:                                        # selectNextVid: tid 0 stays 0
                vid, new = 0, False
            elif tid in active:
                vid, new = active[tid], False
            else:
                vid, new = next_vid, True
            if new:                                             # createThreadInfo / registerFork
                next_vid += 1
                active[tid] = vid
            if t == hbt.SWEV_THREAD_DESTRUCTION:
                active.pop(tid, None)                           # registerExit
            if t == hbt.SWEV_COMM:
                keep[i] = False
            else:
                ev["tag"][i] = vid
        out.append(ev[keep])
    return out


def combinator_merge(streams):
    """Test-side restatement of the reference's binary tree of stream::Combinator nodes
    (hbt/src/tagstack/Stream.h:114-248, makeBinaryTree :190-230) draining `streams` (one
    SWITCH_EVENT_DTYPE array per compute unit) with stop_ts = infinity: every node caches one
    event per input and hands out `a.happenedBefore(b) ? a : b` (Event.h:205-213); an
    exhausted input counts as (kInvalidTimeStamp, max comp_unit_id)."""
    INV = (0xFFFFFFFFFFFFFFFF, 0xFFFF)

    def leaf(arr):
        for i in range(len(arr)):
            yield arr[i]

    def before(a, b):
        ka = (int(a["tstamp"]), int(a["comp_unit_id"])) if a is not None else INV
        kb = (int(b["tstamp"]), int(b["comp_unit_id"])) if b is not None else INV
        if ka[0] > kb[0]:
            return False
        return ka[0] < kb[0] or ka[1] < kb[1]

    def node(ga, gb):
        ea = next(ga, None)
        eb = next(gb, None) if gb is not None else None
        while ea is not None or eb is not None:
            if before(ea, eb):
                yield ea
                ea = next(ga, None)
            else:
                yield eb
                eb = next(gb, None)

    level = [leaf(s) for s in streams]
    while True:
        nxt = [node(level[i], level[i + 1] if i + 1 < len(level) else None)
               for i in range(0, len(level), 2)]
        if len(nxt) == 1:
            out = list(nxt[0])
            res = np.zeros(len(out), dtype=hbt.SWITCH_EVENT_DTYPE)
            for i, e in enumerate(out):
                res[i] = e
            return res
        level = nxt


def counts_outputs(I, E, R):
    return dict(sums=np.zeros((I, 3 + E), np.uint64), scaled=np.zeros((I, E), np.uint64),
                rates=np.zeros((I, E), np.float32), ratios=np.zeros((I, max(R, 1)), np.float64),
                deltas=np.zeros((I, E), np.uint64))


def port_counts(reads: np.ndarray, ratios=()):
    I, Cc, W = reads.shape
    E, R = W - 3, len(ratios)
    o = counts_outputs(I, E, R)
    r = hbt.make_ratios(ratios)
    rc = port().hbt_oracle_counts(_np(np.ascontiguousarray(reads)), I, Cc, E,
                                  C.cast(r, C.c_void_p), R, _np(o["sums"]), _np(o["scaled"]),
                                  _np(o["rates"]), _np(o["ratios"]) if R else None,
                                  _np(o["deltas"]))
    assert rc == 0
    o["ratios"] = o["ratios"][:, :R]
    return o


def counts(reads: np.ndarray, ratios=()):
    """Counter-group oracle for the GPU tests: the C restatement's outputs, cross-checked here
    against the reference's own PerCpuBase::read / GroupReadValues::getCount (oracle/_ref) when
    that library is present (words of a failed-read interval other than word0 are unspecified)."""
    p = port_counts(reads, ratios)
    if have_ref():
        r = ref_counts(reads, ratios)
        ok = r["sums"][:, 0] != 0
        assert np.array_equal(r["sums"][:, 0], p["sums"][:, 0])
        assert np.array_equal(r["sums"][ok], p["sums"][ok])
        for k in ("scaled", "deltas"):
            assert np.array_equal(r[k], p[k]), ("oracle port != reference", k)
        assert np.array_equal(r["rates"], p["rates"], equal_nan=True)
        assert np.array_equal(r["ratios"], p["ratios"], equal_nan=True)
    return p


def ref_counts(reads: np.ndarray, ratios=(), threads=1):
    I, Cc, W = reads.shape
    E, R = W - 3, len(ratios)
    o = counts_outputs(I, E, R)
    r = hbt.make_ratios(ratios)
    t = ref().hbtref_counts(_np(np.ascontiguousarray(reads)), I, Cc, E, C.cast(r, C.c_void_p), R,
                            _np(o["sums"]), _np(o["scaled"]), _np(o["rates"]),
                            _np(o["ratios"]) if R else None, _np(o["deltas"]), threads)
    assert t >= 0, t
    o["ratios"] = o["ratios"][:, :R]
    o["seconds"] = t
    return o


# ---------------------------------------------------------------------------
# Metric tables (dynolog/src/metric_frame): the reference's own classes through oracle/_ref, and a
# numpy / plain-Python restatement for machines without it.
# ---------------------------------------------------------------------------
def ref_frame_query(rows, is_double, frame_len, interval_us, t_last_ns, queries, period_ns, percentage):
    """The REAL MetricFrameVector / MetricFrameTsUnitFixInterval / MetricSeriesSlice.  rows [n][S]
    raw 8-byte values, queries = [(start_ns, end_ns, start_policy, end_policy)].  Returns (length,
    ranges (None where slice() is nullopt), dict of [Q][S] arrays)."""
    lib = ref()
    rows = np.ascontiguousarray(rows).view(np.uint64)
    n, S = rows.shape
    isd = np.ascontiguousarray(is_double, dtype=np.uint8)
    Q = len(queries)
    qs = np.array([q[0] for q in queries], np.int64)
    qe = np.array([q[1] for q in queries], np.int64)
    ps = np.array([q[2] for q in queries], np.int32)
    pe = np.array([q[3] for q in queries], np.int32)
    ranges = np.zeros(Q, dtype=hbt.FRAME_RANGE_DTYPE)
    ok = np.zeros(Q, np.uint8)
    out = dict(diff=np.zeros((Q, S), np.uint64), rate_f64=np.zeros((Q, S), np.float64),
               rate_i64=np.zeros((Q, S), np.int64), avg_f64=np.zeros((Q, S), np.float64),
               avg_i64=np.zeros((Q, S), np.int64), pct=np.zeros((Q, S), np.uint64))
    length = C.c_uint64()
    lib.hbtref_frame_query.restype = C.c_int
    lib.hbtref_frame_query.argtypes = [_P, C.c_uint64, C.c_uint32, _P, C.c_uint32, C.c_int64, C.c_int64, _P, _P, _P,
                                       _P, C.c_uint32, C.c_int64, C.c_float] + [_P] * 8 + [C.POINTER(C.c_uint64)]
    rc = lib.hbtref_frame_query(_np(rows), n, S, _np(isd), frame_len, interval_us, t_last_ns, _np(qs), _np(qe),
                                _np(ps), _np(pe), Q, period_ns, percentage, _np(ranges), _np(ok),
                                _np(out["diff"]), _np(out["rate_f64"]), _np(out["rate_i64"]), _np(out["avg_f64"]),
                                _np(out["avg_i64"]), _np(out["pct"]), C.byref(length))
    assert rc == 0, rc
    rl = [tuple(int(x) for x in ranges[i]) if ok[i] else None for i in range(Q)]
    return int(length.value), rl, out


def _trunc_div(a: int, b: int) -> int:
    """C++ integer division (truncates toward zero)."""
    q = abs(a) // abs(b)
    return q if (a < 0) == (b < 0) else -q


def _wrap64(x: int) -> int:
    x &= (1 << 64) - 1
    return x - (1 << 64) if x >= (1 << 63) else x


def port_frame_range(n_samples, last_time_ns, interval_us, t_start, t_end, pol_s, pol_e):
    """MetricFrameTsUnitFixInterval::getRange / findMatchingOffset (MetricFrameTsUnit.cpp:49-132)."""
    import math

    def match(t, pol):
        if n_samples == 0:
            return None
        off = (n_samples - 1) + float(_trunc_div(t - last_time_ns, 1000)) / float(interval_us)
        top = n_samples - 1
        if pol == hbt.MATCH_CLOSEST:
            o = 0 if off < 0 else top if off > top else int(math.floor(abs(off) + 0.5) * (1 if off >= 0 else -1))
        elif pol == hbt.MATCH_PREV_CLOSEST:
            if off < 0:
                return None
            o = top if off > top else int(math.floor(off))
        else:
            if off > top:
                return None
            o = 0 if off < 0 else int(math.ceil(off))
        return o, last_time_ns - (n_samples - o - 1) * interval_us * 1000
    a, b = match(t_start, pol_s), match(t_end, pol_e)
    if a is None or b is None:
        return None
    return (a[0], b[0], a[1], b[1])


def port_frame_query(rows, is_double, frame_len, interval_us, t_last_ns, queries, period_ns, percentage):
    """Restatement of MetricFrameVector::addSamples (MetricFrame.cpp:82-95), MetricSeries::addSample
    (MetricSeries.h:148-158: the ring keeps the last frame_len samples) and MetricSeriesSlice's
    diff / rate / avg / percentile (MetricFrameBase.h:66-91, MetricSeries.h:189-249)."""
    import struct
    rows = np.ascontiguousarray(rows).view(np.uint64)
    n, S = rows.shape
    held = rows[max(0, n - frame_len):]
    length = len(held)
    Q = len(queries)
    out = dict(diff=np.zeros((Q, S), np.uint64), rate_f64=np.zeros((Q, S), np.float64),
               rate_i64=np.zeros((Q, S), np.int64), avg_f64=np.zeros((Q, S), np.float64),
               avg_i64=np.zeros((Q, S), np.int64), pct=np.zeros((Q, S), np.uint64))
    ranges = []
    f32 = np.float32
    for q, (t0, t1, p0, p1) in enumerate(queries):
        r = port_frame_range(length, t_last_ns, interval_us, t0, t1, p0, p1)
        if r is not None and r[1] < r[0]:
            r = None
        ranges.append(r)
        if r is None:
            continue
        b, e = r[0], r[1] + 1
        duration = r[3] - r[2]
        cnt = e - b
        # lround(float(percentage) * float(n - 1)): round half away from zero, in float
        prod = float(f32(percentage) * f32(cnt - 1))
        nth = int(np.floor(prod + 0.5))
        for s in range(S):
            col = held[b:e, s]
            if is_double[s]:
                v = col.view(np.float64)
                d = float(v[-1] - v[0])
                out["diff"][q, s] = struct.unpack("<Q", struct.pack("<d", d))[0]
                with np.errstate(divide="ignore", invalid="ignore"):
                    out["rate_f64"][q, s] = np.float64(d) * np.float64(period_ns // 1000) / np.float64(_trunc_div(duration, 1000))
                if duration == 0:
                    ri = 0
                elif period_ns > duration:
                    ri = int(d * float(period_ns // duration))
                else:
                    ri = int(d / float(duration // period_ns))
                out["rate_i64"][q, s] = ri
                acc = 0.0
                for x in v:
                    acc = acc + float(x)
                out["avg_f64"][q, s] = acc / cnt
                out["pct"][q, s] = np.sort(v)[nth].view(np.uint64)
            else:
                v = col.view(np.int64)
                d = _wrap64(int(v[-1]) - int(v[0]))
                out["diff"][q, s] = np.uint64(d & ((1 << 64) - 1))
                with np.errstate(divide="ignore", invalid="ignore"):
                    out["rate_f64"][q, s] = np.float64(d) * np.float64(period_ns // 1000) / np.float64(_trunc_div(duration, 1000))
                if duration == 0:
                    ri = 0
                elif period_ns > duration:
                    ri = _wrap64(d * (period_ns // duration))
                else:
                    ri = _trunc_div(d, duration // period_ns)
                out["rate_i64"][q, s] = ri
                acc, acci = 0.0, 0
                for x in v:
                    acc = acc + float(int(x))
                    acci = _wrap64(acci + int(x))
                out["avg_f64"][q, s] = acc / cnt
                out["avg_i64"][q, s] = _trunc_div(acci, cnt)
                out["pct"][q, s] = np.sort(v)[nth].view(np.uint64)
    return length, ranges, out


def frame_query(rows, is_double, frame_len, interval_us, t_last_ns, queries, period_ns, percentage):
    """Oracle for the metric-table tests: the reference's own classes when oracle/_ref is present
    (the restatement is then cross-checked against them), else the restatement."""
    p = port_frame_query(rows, is_double, frame_len, interval_us, t_last_ns, queries, period_ns, percentage)
    if not have_ref():
        return p
    r = ref_frame_query(rows, is_double, frame_len, interval_us, t_last_ns, queries, period_ns, percentage)
    assert r[0] == p[0] and r[1] == p[1], ("oracle port != reference (length / ranges)", r[:2], p[:2])
    live = np.array([x is not None for x in r[1]], dtype=bool)
    for k in r[2]:
        a, b = r[2][k][live], p[2][k][live]
        assert np.array_equal(a.view(np.uint64), b.view(np.uint64)), ("oracle port != reference", k)
    return r


# ---------------------------------------------------------------------------
# Slices.  hbt/src/tagstack needs Boost.Intrusive and hbt/src/mon/MonData.h
# the absent hbt/src/phase, so the reference's own code does not compile here: the checker is a
# plain-Python restatement, pinned to the reference's IntervalSlicerTest.cpp known answers
# (tests/test_slices.py) for the join; the frequencies and the bin matrix have no reference test.
# ---------------------------------------------------------------------------
class ReferenceThrows(Exception):
    """The reference throws on this input (HBT_THROW_ASSERT*)."""


def port_interval_slices(slices, offset, tstamp, duration, trim, abort_on_missing):
    """IntervalSlicer::makeIntervalIntersectingSlices (IntervalSlicer.cpp:52-188) over a list of
    (tstamp, duration, stack_id, swin, swout) tuples; returns (slices or None, num_consumed)."""
    assert duration > 0
    t_end = tstamp + duration
    it = iter(range(offset, len(slices)))
    i = next(it, None)
    consumed = 0
    while i is not None and slices[i][0] + slices[i][1] < tstamp:   # :79-84
        i = next(it, None)
        consumed += 1
    if i is None:
        return None, 0
    if abort_on_missing and slices[i][0] > tstamp:
        return None, 0
    total = 0
    last_end = slices[i][0]
    part = []
    while i is not None and slices[i][0] < t_end:                    # :104-123
        s = slices[i]
        if abort_on_missing and last_end != s[0]:
            return None, 0
        if not trim and s[0] + s[1] > t_end:
            break
        total += s[1]
        part.append(list(s))
        last_end = s[0] + s[1]
        i = next(it, None)
        consumed += 1
    if not part:
        return None, 0
    last_slice_end = part[-1][0] + part[-1][1]
    if abort_on_missing and last_slice_end < t_end:
        return None, 0
    if trim:                                                         # :147-161
        f = part[0]
        if f[0] < tstamp:
            overlap = f[0] + f[1] - tstamp
            total -= f[1] - overlap
            f[1] = overlap
            f[0] = tstamp
            f[3] = hbt.TT_ANALYSIS
    if last_slice_end > t_end:
        consumed -= 1
    if trim and last_slice_end > t_end:
        delta = last_slice_end - t_end
        part[-1][1] -= delta
        part[-1][4] = hbt.TT_ANALYSIS
        total -= delta
    if total > duration:   # HBT_THROW_ASSERT_IF(slices_duration > duration) (:179-181)
        raise ReferenceThrows("slices_duration > duration")
    return [tuple(x) for x in part], consumed


def port_consume_slices(slices, intervals, trim, abort_on_missing, offset=0):
    """consumeSlices (:196-216) once per interval with the consumer mark carried along."""
    out = []
    for (t, d) in intervals:
        part, n = port_interval_slices(slices, offset, t, d, trim, abort_on_missing)
        out.append(part or [])
        offset += n
    return out, offset


def port_stack_id_freqs(per_interval_slices):
    """CuMonData::stackIdFreqs (MonData.h:245-266): {stack_id: [duration, num_obs, num_intervals]}."""
    freqs = {}
    for sl in per_interval_slices:
        seen = set()
        for s in sl:
            f = freqs.setdefault(s[2], [0, 0, 0])
            f[1] += 1
            f[0] = (f[0] + s[1]) & ((1 << 64) - 1)
            if s[2] not in seen:
                f[2] += 1
                seen.add(s[2])
    return freqs


def port_accum_freqs(per_unit_freqs):
    """MonData::stackIdFreqs (:386-403) with SliceFreq::accum (:41-45)."""
    tot = {}
    for fs in per_unit_freqs:
        for sid, f in fs.items():
            t = tot.setdefault(sid, [0, 0, 0])
            for i in range(3):
                t[i] = (t[i] + f[i]) & ((1 << 64) - 1)
    return tot


def port_bin_matrix(per_unit_pis, freqs, min_distinct_intervals, min_duration, n_rows):
    """TagStackIdBinner (MonData.h:517-606) -> {stack_id: column of durations per row}; ignored stack
    ids share the column keyed None.  (Columns are keyed by stack id because the reference's bin
    numbering follows an unordered_map's iteration order.)"""
    keep = {sid for sid, f in freqs.items() if f[2] >= min_distinct_intervals and f[0] >= min_duration}
    cols = {sid: [0] * n_rows for sid in keep}
    cols[None] = [0] * n_rows
    for pis in per_unit_pis:
        i = 0
        for sl in pis:
            if not sl:
                continue
            row_dur = 0
            for s in sl:
                if s[1] == 0:
                    continue
                row_dur += s[1]
                col = cols[s[2]] if s[2] in keep else cols[None]
                col[i] = (col[i] + s[1]) & ((1 << 64) - 1)
            if row_dur:
                i += 1
    return cols
