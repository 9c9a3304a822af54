"""Synthetic perf_event traces: ring images of PERF_RECORD_*
streams and recorded counter-group reads, generated with torch so the same code
runs on the CPU (tests) and on the GPU (bench.py at full size).

Everything is a pure function of a 64-bit seed through splitmix64, so a trace is
identical on every device.  Streams are int64 tensors of little-endian u64 words
(records are u64 aligned, /usr/include/linux/perf_event.h:815-819).
"""
from __future__ import annotations

import torch

REC_MMAP, REC_LOST, REC_COMM, REC_EXIT, REC_THROTTLE, REC_UNTHROTTLE, REC_FORK = 1, 2, 3, 4, 5, 6, 7
REC_READ, REC_SAMPLE, REC_AUX, REC_ITRACE_START, REC_SWITCH_CPU_WIDE = 8, 9, 11, 12, 15

_M64 = (1 << 64) - 1


def _s64(x: int) -> int:
    """Python int -> the int64 with the same low 64 bits."""
    x &= _M64
    return x - (1 << 64) if x >= (1 << 63) else x


def stream_seed(config: int, cpu: int) -> int:
    """stream seed = 0x5EED000000000000 + (config# << 32) + cpu."""
    return 0x5EED000000000000 + (config << 32) + cpu


def _lsr(x: torch.Tensor, k: int) -> torch.Tensor:
    return (x >> k) & ((1 << (64 - k)) - 1)


def splitmix64(x: torch.Tensor) -> torch.Tensor:
    z = x + _s64(0x9E3779B97F4A7C15)
    z = (z ^ _lsr(z, 30)) * _s64(0xBF58476D1CE4E5B9)
    z = (z ^ _lsr(z, 27)) * _s64(0x94D049BB133111EB)
    return z ^ _lsr(z, 31)


def rand_u64(seed: int, n: int, device="cpu", stream: int = 0) -> torch.Tensor:
    """n pseudo-random 64-bit words (as int64)."""
    idx = torch.arange(n, dtype=torch.int64, device=device)
    base = _s64(seed * 0x2545F4914F6CDD1D + stream * 0x9E3779B97F4A7C15)
    return splitmix64(splitmix64(idx + base))


def rand_uniform(seed: int, n: int, device="cpu", stream: int = 0) -> torch.Tensor:
    """n doubles in [0, 1)."""
    return _lsr(rand_u64(seed, n, device, stream), 11).to(torch.float64) * (1.0 / (1 << 53))


def rand_below(seed: int, n: int, bound: int, device="cpu", stream: int = 0) -> torch.Tensor:
    return (rand_uniform(seed, n, device, stream) * bound).to(torch.int64).clamp_(max=bound - 1)


def zipf_cdf(n: int, s: float, device="cpu") -> torch.Tensor:
    w = torch.arange(1, n + 1, dtype=torch.float64, device=device).pow(-s)
    c = torch.cumsum(w, 0)
    return c / c[-1]


def zipf_sample(seed: int, n: int, cdf: torch.Tensor, stream: int = 0) -> torch.Tensor:
    u = rand_uniform(seed, n, cdf.device, stream)
    return torch.searchsorted(cdf, u).clamp_(max=cdf.numel() - 1)


def header(rtype: int, size_bytes, misc: int = 0):
    """perf_event_header {u32 type; u16 misc; u16 size} as one u64 word."""
    return rtype | (misc << 32) | (size_bytes << 48)


# ---------------------------------------------------------------------------
# Streams
# ---------------------------------------------------------------------------
def group_sample_stream(seed: int, n_samples: int, n_events: int, device="cpu", lost_every: int = 0,
                        t0: int = 1_000_000, regress_at: int = -1) -> torch.Tensor:
    """mode::Sampling stream: SAMPLE = hdr tstamp nr time_enabled time_running
    value[nr] (40+8nr B); with lost_every > 0 a Sampling::Lost (hdr id num_lost,
    24 B) precedes every lost_every-th sample.  Values are cumulative counters,
    tstamps strictly increase (regress_at >= 0 makes that sample's tstamp go
    backwards)."""
    E = n_events
    W = 5 + E
    n = n_samples
    dt = 1000 + rand_below(seed, n, 9000, device, 1)
    ts = t0 + torch.cumsum(dt, 0)
    if regress_at >= 0:
        ts[regress_at] = ts[max(regress_at - 1, 0)] - 5
    inc = 100_000 + rand_below(seed, n * E, 9_900_000, device, 2).view(n, E)
    val = torch.cumsum(inc, 0)
    te = torch.cumsum(dt, 0)
    tr = te - rand_below(seed, n, 7, device, 3)
    rec = torch.empty((n, W), dtype=torch.int64, device=device)
    rec[:, 0] = header(REC_SAMPLE, 8 * W)
    rec[:, 1] = ts
    rec[:, 2] = E
    rec[:, 3] = te
    rec[:, 4] = tr
    rec[:, 5:] = val
    if lost_every <= 0:
        return rec.reshape(-1)
    is_lost = (torch.arange(n, device=device) % lost_every) == (lost_every - 1)
    extra = is_lost.to(torch.int64) * 3
    off = torch.cumsum(W + extra, 0) - (W + extra)  # start of each (lost?)+sample group
    total = int((W * n + extra.sum()).item())
    out = torch.zeros(total, dtype=torch.int64, device=device)
    s_off = off + extra
    out[(s_off[:, None] + torch.arange(W, device=device)[None, :]).reshape(-1)] = rec.reshape(-1)
    l_off = off[is_lost]
    nl = l_off.numel()
    out[l_off] = header(REC_LOST, 24)
    out[l_off + 1] = 7
    out[l_off + 2] = 1 + rand_below(seed, nl, 50, device, 4)
    return out


def ip_sample_stream(seed: int, n_samples: int, cpu: int, device="cpu", n_pids: int = 4096,
                     n_ips: int = 65536, s_pid: float = 1.1, s_ip: float = 1.1,
                     sample_words: int = 5, cdfs=None, t0: int = 1_000_000,
                     lost_every: int = 0) -> torch.Tensor:
    """sample_type = IP|TID|TIME|CPU samples (hdr ip pid,tid time cpu,res [extra]):
    pid ~ Zipf(s_pid) over n_pids, ip ~ Zipf(s_ip) over n_ips code addresses per
    pid; time strictly increasing."""
    n = n_samples
    if cdfs is None:
        cdfs = (zipf_cdf(n_pids, s_pid, device), zipf_cdf(n_ips, s_ip, device))
    pid_idx = zipf_sample(seed, n, cdfs[0], 11)
    ip_idx = zipf_sample(seed, n, cdfs[1], 12)
    pid = 100 + pid_idx * 7
    tid = pid + (ip_idx & 3)
    # code address: a per-(pid, ip_idx) pseudo-random canonical kernel/user address
    ip = splitmix64(pid_idx * 65536 + ip_idx + _s64(0xC0DE0000)) | _s64(0xFFFF800000000000)
    ip = torch.where((pid_idx & 1) == 1, ip & _s64(0x00007FFFFFFFFFF0), ip)
    dt = 1000 + rand_below(seed, n, 9000, device, 13)
    ts = t0 + torch.cumsum(dt, 0)
    W = sample_words
    rec = torch.zeros((n, W), dtype=torch.int64, device=device)
    rec[:, 0] = header(REC_SAMPLE, 8 * W)
    rec[:, 1] = ip
    rec[:, 2] = pid | (tid << 32)
    rec[:, 3] = ts
    rec[:, 4] = cpu
    if W > 5:
        rec[:, 5:] = rand_u64(seed, n * (W - 5), device, 14).view(n, W - 5)
    if lost_every <= 0:
        return rec.reshape(-1)
    is_lost = (torch.arange(n, device=device) % lost_every) == (lost_every - 1)
    extra = is_lost.to(torch.int64) * 3
    off = torch.cumsum(W + extra, 0) - (W + extra)
    total = int((W * n + extra.sum()).item())
    out = torch.zeros(total, dtype=torch.int64, device=device)
    s_off = off + extra
    out[(s_off[:, None] + torch.arange(W, device=device)[None, :]).reshape(-1)] = rec.reshape(-1)
    l_off = off[is_lost]
    out[l_off] = header(REC_LOST, 24)
    out[l_off + 1] = 3
    out[l_off + 2] = 1 + rand_below(seed, l_off.numel(), 9, device, 15)
    return out


# Record mix of BASELINE config 4 (by count).
MIXED_MIX = (("sample", 0.70), ("switch", 0.15), ("comm", 0.05), ("mmap", 0.05), ("lost", 0.001),
             ("fork", 0.0245), ("exit", 0.0245))


# Every record type consume() dispatches on (CpuEventsGroup.h:1372-1501): the config-4 mix plus
# THROTTLE / UNTHROTTLE (hdr time id stream_id), READ (mode::Sampling::Read, :247-259), AUX and
# ITRACE_START (mode::AUXSpace, :205-227).
FULL_MIX = (("sample", 0.55), ("switch", 0.15), ("comm", 0.05), ("mmap", 0.05), ("lost", 0.01),
            ("fork", 0.03), ("exit", 0.03), ("throttle", 0.02), ("unthrottle", 0.02), ("read", 0.04),
            ("aux", 0.03), ("itrace", 0.02))
_TRAILER_KINDS = ("switch", "comm", "mmap", "lost", "fork", "exit")


def mixed_stream(seed: int, n_records: int, n_events: int, device="cpu", mix=MIXED_MIX,
                 t0: int = 1_000_000, unknown_types: bool = True) -> torch.Tensor:
    """Mixed stream (config 4) in the harness-local Mixed mode: ContextSwitch
    records with the {pid,tid,time} sample_id trailer (CpuEventsGroup.h:105-184),
    Sampling SAMPLE records (:268-281), and MMAP records that no reference mode
    handles (they take consume()'s default: branch, :1498-1500)."""
    n = n_records
    u = rand_uniform(seed, n, device, 21)
    kinds = [k for k, _ in mix]
    edges = torch.tensor([p for _, p in mix], dtype=torch.float64, device=device).cumsum(0)
    edges = edges / edges[-1]
    kind = torch.searchsorted(edges, u).clamp_(max=len(kinds) - 1)
    E = n_events
    # size in words per record
    name_w = 1 + rand_below(seed, n, 2, device, 22)     # comm: 1..16 B name, padded to 8
    file_w = 1 + rand_below(seed, n, 32, device, 23)    # mmap: 8..256 B filename
    words = torch.zeros(n, dtype=torch.int64, device=device)
    sizes = {"sample": 5 + E, "switch": 4, "lost": 5, "fork": 6, "exit": 6, "throttle": 4,
             "unthrottle": 4, "read": 5 + E, "aux": 4, "itrace": 2}
    for i, k in enumerate(kinds):
        sel = kind == i
        if k == "comm":
            words = torch.where(sel, 2 + name_w + 2, words)
        elif k == "mmap":
            words = torch.where(sel, 5 + file_w + 2, words)
        else:
            words = torch.where(sel, torch.full_like(words, sizes[k]), words)
    off = torch.cumsum(words, 0) - words
    total = int(words.sum().item())
    out = rand_u64(seed, total, device, 24) & _s64(0x0000FFFFFFFFFFFF)  # payload noise, top 16 bits 0
    dt = 1000 + rand_below(seed, n, 9000, device, 25)
    ts = t0 + torch.cumsum(dt, 0)
    pid = 100 + rand_below(seed, n, 4096, device, 26)
    pidtid = pid | ((pid + 1) << 32)
    types = {"sample": REC_SAMPLE, "switch": REC_SWITCH_CPU_WIDE, "comm": REC_COMM,
             "mmap": REC_MMAP, "lost": REC_LOST, "fork": REC_FORK, "exit": REC_EXIT,
             "throttle": REC_THROTTLE, "unthrottle": REC_UNTHROTTLE, "read": REC_READ,
             "aux": REC_AUX, "itrace": REC_ITRACE_START}
    hdr = torch.zeros(n, dtype=torch.int64, device=device)
    for i, k in enumerate(kinds):
        hdr = torch.where(kind == i, types[k] | (words << 51), hdr)  # size = 8*words -> << 48+3
    if unknown_types:
        # a few MMAP records become type 10 (MMAP2) and 19 (CGROUP): also unhandled
        alt = rand_below(seed, n, 8, device, 27)
        is_mmap = kind == kinds.index("mmap")
        hdr = torch.where(is_mmap & (alt == 0), (hdr & ~0xFFFFFFFF) | 10, hdr)
        hdr = torch.where(is_mmap & (alt == 1), (hdr & ~0xFFFFFFFF) | 19, hdr)
    out[off] = hdr
    # sample_id trailer {pid,tid ; time} on the context-switch style records
    non_sample = torch.zeros(n, dtype=torch.bool, device=device)
    for i, k in enumerate(kinds):
        if k in _TRAILER_KINDS:
            non_sample |= kind == i
    o = off[non_sample]
    w = words[non_sample]
    out[o + w - 2] = pidtid[non_sample]
    out[o + w - 1] = ts[non_sample]
    out[o + 1] = pidtid[non_sample]
    # SAMPLE bodies: cumulative values
    sel = kind == kinds.index("sample")
    so = off[sel]
    ns = so.numel()
    if ns:
        inc = 100_000 + rand_below(seed, ns * E, 9_900_000, device, 28).view(ns, E)
        val = torch.cumsum(inc, 0)
        out[so + 1] = ts[sel]
        out[so + 2] = E
        out[so + 3] = ts[sel] - t0
        out[so + 4] = ts[sel] - t0
        out[(so[:, None] + 5 + torch.arange(E, device=device)[None, :]).reshape(-1)] = val.reshape(-1)
    # LOST: hdr id num_lost sample_id
    lo = off[kind == kinds.index("lost")]
    if lo.numel():
        out[lo + 1] = 5
        out[lo + 2] = 1 + rand_below(seed, lo.numel(), 100, device, 29)
    if "read" in kinds:   # hdr {pid,tid} nr time_enabled time_running value[nr]
        sel = kind == kinds.index("read")
        ro = off[sel]
        if ro.numel():
            out[ro + 1] = pidtid[sel]
            out[ro + 2] = E
            out[ro + 3] = ts[sel] - t0
            out[ro + 4] = ts[sel] - t0
    for k in ("throttle", "unthrottle"):   # hdr time id stream_id
        if k in kinds:
            sel = kind == kinds.index(k)
            out[off[sel] + 1] = ts[sel]
    if "itrace" in kinds:  # hdr {pid,tid}
        sel = kind == kinds.index("itrace")
        out[off[sel] + 1] = pidtid[sel]
    return out


MISC_SWITCH_OUT = 1 << 13          # PERF_RECORD_MISC_SWITCH_OUT
MISC_SWITCH_OUT_PREEMPT = 1 << 14  # PERF_RECORD_MISC_SWITCH_OUT_PREEMPT

CTX_MIX = (("switch", 0.86), ("comm", 0.04), ("fork", 0.04), ("exit", 0.04), ("lost", 0.02))


def ctx_switch_stream(seed: int, n_records: int, device="cpu", mix=CTX_MIX, t0: int = 1_000_000,
                      n_tids: int = 512, fork_tid_base: int = 0) -> torch.Tensor:
    """A mode::ContextSwitch stream (what CpuThreadSwitchGenerator's sw_dummy event
    produces, CpuEventsGroup.h:105-184): SWITCH_CPU_WIDE in / out-yield / out-preempt,
    COMM (8 or 16 byte names), FORK, EXIT and LOST records with {pid,tid,time} sample_id
    trailers and increasing timestamps.  Vectorised.  Switch / COMM / EXIT records draw their
    tid from a pool of n_tids; every FORK creates a fresh tid above the pool (the reference's
    handlers assert on a second FORK of a live thread), so the reference's own generator
    accepts the stream, although threads "exit" without having forked."""
    n = n_records
    u = rand_uniform(seed, n, device, 41)
    kinds = [k for k, _ in mix]
    edges = torch.tensor([p for _, p in mix], dtype=torch.float64, device=device).cumsum(0)
    edges = edges / edges[-1]
    kind = torch.searchsorted(edges, u).clamp_(max=len(kinds) - 1)
    name_w = 1 + rand_below(seed, n, 2, device, 42)
    base_w = {"switch": 4, "lost": 5, "fork": 6, "exit": 6}
    types = {"switch": REC_SWITCH_CPU_WIDE, "comm": REC_COMM, "lost": REC_LOST, "fork": REC_FORK,
             "exit": REC_EXIT}
    words = torch.zeros(n, dtype=torch.int64, device=device)
    rtype = torch.zeros(n, dtype=torch.int64, device=device)
    for i, k in enumerate(kinds):
        sel = kind == i
        w = 2 + name_w + 2 if k == "comm" else torch.full_like(words, base_w[k])
        words = torch.where(sel, w, words)
        rtype = torch.where(sel, torch.full_like(rtype, types[k]), rtype)
    off = torch.cumsum(words, 0) - words
    total = int(words.sum().item())
    out = rand_u64(seed, total, device, 43) & _s64(0x0000FFFFFFFFFFFF)
    dt = 1000 + rand_below(seed, n, 9000, device, 44)
    ts = t0 + torch.cumsum(dt, 0)
    tid = rand_below(seed, n, n_tids, device, 45)            # tid 0 (the idle task) occurs too ...
    # ... but only in SWITCH records: the reference's COMM / EXIT handlers look tid 0 up in a map
    # that never holds it (PerCpuThreadSwitchGenerator.h:92-97, 279, 128) and abort
    tid = torch.where((rtype != REC_SWITCH_CPU_WIDE) & (tid == 0), torch.ones_like(tid), tid)
    pid = tid - (tid % 4)
    pidtid = pid | (tid << 32)
    # SWITCH misc: in / out-yield / out-preempt
    sw = rand_below(seed, n, 3, device, 46)
    misc = torch.where(sw == 0, torch.zeros_like(sw),
                       torch.where(sw == 1, torch.full_like(sw, MISC_SWITCH_OUT),
                                   torch.full_like(sw, MISC_SWITCH_OUT | MISC_SWITCH_OUT_PREEMPT)))
    misc = torch.where(rtype == REC_SWITCH_CPU_WIDE, misc, torch.zeros_like(misc))
    out[off] = rtype | (misc << 32) | (words << 51)
    out[off + words - 2] = pidtid      # sample_id {pid, tid}
    out[off + words - 1] = ts          # sample_id.time
    out[off + 1] = pidtid              # switch: next_prev ; comm: pid,tid ; lost: id
    is_fork = rtype == REC_FORK
    # (fork_tid_base keeps the fresh tids of different CPUs' streams apart: they share one thread map)
    tid = torch.where(is_fork, n_tids + fork_tid_base + torch.cumsum(is_fork.to(torch.int64), 0), tid)
    pid = torch.where(is_fork, tid, pid)
    fe = is_fork | (rtype == REC_EXIT)
    o = off[fe]
    if o.numel():                      # Exit/Fork: pid ppid | tid ptid | time
        out[o + 1] = pid[fe] | ((pid[fe] + 1) << 32)
        out[o + 2] = tid[fe] | ((tid[fe] + 1) << 32)
        out[o + 3] = ts[fe]
    lo = off[rtype == REC_LOST]
    if lo.numel():
        out[lo + 1] = 7
        out[lo + 2] = 1 + rand_below(seed, lo.numel(), 100, device, 47)
    return out


def ctx_switch_stream_consistent(seed: int, n_records: int, t0: int = 1_000_000, lost_every: int = 0,
                                 first_tid: int = 1000) -> torch.Tensor:
    """Small context-switch stream with consistent thread lifetimes, as the reference's
    handlers require (they assert on a second FORK of a live thread or an EXIT before the
    FORK): tids are forked once, switch in and out, get renamed, and exit once.  Pure
    Python; for the reference-pinning tests."""
    import random
    rng = random.Random(seed)
    live, next_tid, ts = [], first_tid, t0
    words = []

    def hdr(rtype, n_words, misc=0):
        return rtype | (misc << 32) | ((8 * n_words) << 48)

    for i in range(n_records):
        ts += rng.randint(1000, 10000)
        r = rng.random()
        if lost_every and i % lost_every == lost_every - 1:
            words += [hdr(REC_LOST, 5), 7, rng.randint(1, 100), 4 | (4 << 32), ts]
        elif r < 0.12 or not live:
            tid = next_tid
            next_tid += 1
            live.append(tid)
            words += [hdr(REC_FORK, 6), (tid - tid % 4) | (1 << 32), tid | (1 << 32), ts,
                      1 | (1 << 32), ts]
        elif r < 0.20 and len(live) > 2:
            tid = live.pop(rng.randrange(len(live)))
            words += [hdr(REC_EXIT, 6), (tid - tid % 4) | (1 << 32), tid | (1 << 32), ts,
                      (tid - tid % 4) | (tid << 32), ts]
        elif r < 0.28:
            tid = rng.choice(live)
            nw = rng.randint(1, 2)
            words += [hdr(REC_COMM, 2 + nw + 2), (tid - tid % 4) | (tid << 32)] + \
                [0x0000616161616161] * nw + [(tid - tid % 4) | (tid << 32), ts]
        else:
            tid = rng.choice(live + [0])
            misc = rng.choice([0, MISC_SWITCH_OUT, MISC_SWITCH_OUT | MISC_SWITCH_OUT_PREEMPT])
            words += [hdr(REC_SWITCH_CPU_WIDE, 4, misc), 5 | (6 << 32), (tid - tid % 4) | (tid << 32), ts]
    return torch.tensor([_s64(w) for w in words], dtype=torch.int64)


# ---------------------------------------------------------------------------
# Ring placement
# ---------------------------------------------------------------------------
def place_in_ring(stream: torch.Tensor, data_size: int, tail: int, fill_seed: int = 0):
    """Lay `stream` (int64 words) into a ring image of data_size bytes starting at
    absolute position `tail` (wrapping); the rest of the ring holds stale noise.
    Returns (ring words int64 [data_size/8], head)."""
    assert data_size & (data_size - 1) == 0 and tail % 8 == 0
    nw = data_size // 8
    n = stream.numel()
    assert n <= nw, "stream larger than the ring"
    dev = stream.device
    if fill_seed:
        ring = rand_u64(fill_seed, nw, dev, 31) & _s64(0x0000FFFFFFFFFFFF)
    else:
        ring = torch.zeros(nw, dtype=torch.int64, device=dev)
    idx = (torch.arange(n, dtype=torch.int64, device=dev) + (tail // 8)) & (nw - 1)
    ring[idx] = stream
    return ring, tail + 8 * n


def random_tail(seed: int, data_size: int, laps: int = 3) -> int:
    """A u64-aligned tail a few laps into the position space, not tile aligned, so
    that the window wraps the ring end."""
    r = int(rand_u64(seed, 1, "cpu", 41).item()) & _M64
    return laps * data_size + (data_size // 2) + (r % (data_size // 4)) // 8 * 8


# ---------------------------------------------------------------------------
# Counter-group traces
# ---------------------------------------------------------------------------
def counts_trace(seed: int, n_intervals: int, n_cpus: int, n_events: int, device="cpu",
                 multiplexed: bool = False) -> torch.Tensor:
    """reads[I][C][3+E] u64 (as int64): one GroupReadValues<TMode>::T
    (CpuEventsGroup.h:371-379) per (interval, cpu), cumulative since enable():
    count += U[1e5,1e7] per interval, time_enabled += 1e9 +- jitter,
    time_running = time_enabled, or (multiplexed) += d_enabled * U[0.25,1]
    (monotone, 0 < running <= enabled)."""
    I, Cc, E = n_intervals, n_cpus, n_events
    inc = 100_000 + rand_below(seed, I * Cc * E, 9_900_000, device, 51).view(I, Cc, E)
    cnt = torch.cumsum(inc, 0)
    dte = 1_000_000_000 - 50_000 + rand_below(seed, I * Cc, 100_000, device, 52).view(I, Cc)
    te = torch.cumsum(dte, 0)
    if multiplexed:
        frac = 0.25 + 0.75 * rand_uniform(seed, I * Cc, device, 53).view(I, Cc)
        dtr = (dte.to(torch.float64) * frac).to(torch.int64).clamp_(min=1)
        tr = torch.cumsum(dtr, 0)
    else:
        tr = te.clone()
    out = torch.empty((I, Cc, 3 + E), dtype=torch.int64, device=device)
    out[:, :, 0] = E
    out[:, :, 1] = te
    out[:, :, 2] = tr
    out[:, :, 3:] = cnt
    return out
