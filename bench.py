#!/usr/bin/env python
"""bench.py -- records decoded+aggregated per second on synthetic perf_event traces.

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload ...]

Default workload = BASELINE.json configs[1]: PERF_RECORD_SAMPLE ring-buffer decode
+ (cpu,pid,ip) histogram, 256 CPUs x 16 MiB rings per GPU (IP|TID|TIME|CPU samples
of 40 B, pid ~ Zipf(1.1) over 4096, ip ~ Zipf(1.1) over 65536 addresses per pid).
One "step" = one pass of decode+aggregate over that batch.  Under torchrun every
rank holds its own 256-CPU shard (weak scaling, sharded by CPU id); the decode needs
no exchange, and each timed step ends with the gather of every rank's per-CPU results
on every rank (hbt_fetch_results_all: one NCCL all-gather inside the timed region).

The line also carries a "diagnostics" block: the other BASELINE configs (C1 on the host
cores, C3 deltas, C4 mixed stream / switch events, C5 counter groups -- at N > 1 sharded by
CPU with the integer all-reduce inside the timed step), each with its own clock sample.

Prints ONE JSON line (rank 0).  `value` is measured with the traces resident in
HBM; `e2e` goes through the host-buffer C-ABI calls with the H2D / D2H copies
inside the timed region.  `--impl reference` times the reference's own consume()
(oracle/_ref, built from the reference headers) on the host cores instead.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

import dynolog_1_b200 as hbt  # noqa: E402
from dynolog_1_b200 import tracegen as tg  # noqa: E402

METRIC = "perf records decoded+aggregated/sec"
UNIT = "records/s"


def load_traffic(kernel: str):
    """dram bytes per launch of `kernel` from the last committed ncu capture, or None."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        with open(p) as f:
            return json.load(f)[kernel]["traffic_bytes"]
    except Exception:
        return None


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons, sampled every 20 ms for the whole life of the bench
    process (started once, long before the first timed region: its start-up takes the driver's
    locks for a moment and would otherwise stall the first timed step of every rank on the box).
    window(t0, t1) reports the samples read inside [t0, t1]; a region shorter than the sampling
    period gets the samples closest to it."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index: int):
        self.index = index
        self.samples = []
        self.proc = None

    def start(self):
        if self.proc is not None:
            return self
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "20"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append((time.time(), line.strip()))

    def window(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"], "samples": 0}
        deadline = time.time() + 0.5
        while not self.samples and time.time() < deadline:   # never return empty-handed
            time.sleep(0.02)
        snap = list(self.samples)
        lines = [l for ts, l in snap if t0 <= ts <= t1 + 0.02]
        nearest = False
        if not lines:
            lines = [l for _, l in sorted(snap, key=lambda x: abs(x[0] - 0.5 * (t0 + t1)))[:3]]
            nearest = True
        sm, mx, reasons = [], [], set()
        for s_ in lines:
            parts = [x.strip() for x in s_.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for n, v in zip(self.NAMES, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        out = {"sm_mhz": statistics.median(sm) if sm else None,
               "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons), "samples": len(sm)}
        if nearest:
            out["note"] = "region shorter than the 20 ms sampling period: nearest samples"
        return out

    def stop(self, t0=None, t1=None):
        """Last use: report [t0, t1] and end the sampling process."""
        out = self.window(t0 if t0 is not None else 0.0, t1 if t1 is not None else time.time())
        self.close()
        return out

    def close(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
            self.proc = None


_SAMPLER = None


def sampler_for(local: int) -> ClockSampler:
    """The process-wide clock sampler (rank 0 only ever asks for it)."""
    global _SAMPLER
    if _SAMPLER is None:
        _SAMPLER = ClockSampler(local).start()
    return _SAMPLER


# ---------------------------------------------------------------------------
# workload: config 2
# ---------------------------------------------------------------------------
def make_c2(n_cpus: int, ring_bytes: int, cpu_base: int, device, sample_words=5):
    """n_cpus ring images on `device`; returns (imgs[int64 n_cpus x words], tails, heads, n_records)."""
    imgs = torch.empty((n_cpus, ring_bytes // 8), dtype=torch.int64, device=device)
    tails, heads = [], []
    n = (ring_bytes - 65536) // (8 * sample_words)  # slack: the ring is not completely full
    cdfs = (tg.zipf_cdf(4096, 1.1, device), tg.zipf_cdf(65536, 1.1, device))
    for c in range(n_cpus):
        cpu = cpu_base + c
        seed = tg.stream_seed(2, cpu)
        s = tg.ip_sample_stream(seed, n, cpu=cpu, device=device, sample_words=sample_words, cdfs=cdfs)
        tail = tg.random_tail(seed, ring_bytes)  # not tile aligned: >= 1 record wraps the ring end
        ring, head = tg.place_in_ring(s, ring_bytes, tail, fill_seed=seed ^ 0xABCDEF)
        imgs[c] = ring
        tails.append(tail)
        heads.append(head)
    return imgs, tails, heads, n * n_cpus


def make_group_rings(n_cpus: int, ring_bytes: int, cpu_base: int, device, n_events: int,
                     lost_every: int = 5000):
    """Rings of mode::Sampling SAMPLE records (40 + 8*n_events bytes) with a LOST now and then."""
    imgs = torch.empty((n_cpus, ring_bytes // 8), dtype=torch.int64, device=device)
    tails, heads = [], []
    rec = 40 + 8 * n_events
    n = (ring_bytes - 65536) // (rec + 1)
    total = 0
    for c in range(n_cpus):
        seed = tg.stream_seed(3, cpu_base + c)
        s = tg.group_sample_stream(seed, n, n_events, device=device, lost_every=lost_every)
        tail = tg.random_tail(seed, ring_bytes)
        ring, head = tg.place_in_ring(s, ring_bytes, tail, fill_seed=seed ^ 0xABCDEF)
        imgs[c] = ring
        tails.append(tail)
        heads.append(head)
        total += n + n // lost_every
    return imgs, tails, heads, total


def make_hostile_rings(n_cpus: int, ring_bytes: int, cpu_base: int, device, n_events: int):
    """Group SAMPLE rings whose counter VALUES are valid-looking perf_event_headers (a SAMPLE header
    of the stream's own size in value[2], a LOST header in value[4]): chains of fake records that even
    validate, so segment-entry guesses that land on them go wrong and have to be caught and re-run."""
    imgs = torch.empty((n_cpus, ring_bytes // 8), dtype=torch.int64, device=device)
    tails, heads = [], []
    W = 5 + n_events
    n = (ring_bytes - 65536) // (8 * W)
    for c in range(n_cpus):
        seed = tg.stream_seed(7, cpu_base + c)
        s = tg.group_sample_stream(seed, n, n_events, device=device)
        s[7::W] = tg.header(tg.REC_SAMPLE, 8 * W)
        s[9::W] = tg.header(tg.REC_LOST, 24)
        tail = tg.random_tail(seed, ring_bytes)
        ring, head = tg.place_in_ring(s, ring_bytes, tail, fill_seed=0)
        imgs[c] = ring
        tails.append(tail)
        heads.append(head)
    return imgs, tails, heads, n * n_cpus


def make_mixed_rings(n_cpus: int, ring_bytes: int, cpu_base: int, device, n_events: int):
    """Rings of the mixed stream of config 4 (SAMPLE 70 % / SWITCH_CPU_WIDE 15 % / COMM / MMAP /
    LOST / FORK / EXIT, 24-320 B records)."""
    imgs = torch.empty((n_cpus, ring_bytes // 8), dtype=torch.int64, device=device)
    tails, heads = [], []
    avg = 0.70 * (40 + 8 * n_events) + 0.15 * 32 + 0.05 * 44 + 0.05 * 188 + 0.001 * 40 + 0.049 * 48
    n = int((ring_bytes - (1 << 20)) / (avg * 1.02))
    for c in range(n_cpus):
        seed = tg.stream_seed(4, cpu_base + c)
        s = tg.mixed_stream(seed, n, n_events, device=device)
        assert s.numel() * 8 <= ring_bytes - 4096, "mixed stream does not fit the ring"
        tail = tg.random_tail(seed, ring_bytes)
        ring, head = tg.place_in_ring(s, ring_bytes, tail, fill_seed=seed ^ 0xABCDEF)
        imgs[c] = ring
        tails.append(tail)
        heads.append(head)
        del s, ring
    return imgs, tails, heads, n * n_cpus


def make_switch_rings(n_cpus: int, ring_bytes: int, cpu_base: int, device):
    """Rings of mode::ContextSwitch records (what CpuThreadSwitchGenerator's sw_dummy event
    fills its buffer with): SWITCH_CPU_WIDE 86 % / COMM / FORK / EXIT / LOST, 32-48 B records."""
    imgs = torch.empty((n_cpus, ring_bytes // 8), dtype=torch.int64, device=device)
    tails, heads = [], []
    n = int((ring_bytes - (1 << 20)) / (34.4 * 1.02))
    for c in range(n_cpus):
        seed = tg.stream_seed(6, cpu_base + c)
        s = tg.ctx_switch_stream(seed, n, device=device, fork_tid_base=(cpu_base + c) << 22)
        assert s.numel() * 8 <= ring_bytes - 4096, "switch stream does not fit the ring"
        tail = tg.random_tail(seed, ring_bytes)
        ring, head = tg.place_in_ring(s, ring_bytes, tail, fill_seed=seed ^ 0xABCDEF)
        imgs[c] = ring
        tails.append(tail)
        heads.append(head)
        del s, ring
    return imgs, tails, heads, n * n_cpus


class _DevArray:
    """A device buffer of the library seen as a torch tensor (__cuda_array_interface__)."""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (ptr, False), "version": 2}


def hist_weight_per_cpu(ctx, n_cpus: int, dev) -> np.ndarray:
    """Sum of the entry weights of every CPU's lists, computed on the device straight from the
    list storage (hbt_hist_lists_dev): [n_cpus] int64."""
    ent_p, cur_p, lpc, cap = ctx.hist_lists_dev()
    out = np.zeros(n_cpus, dtype=np.int64)
    cursors = torch.as_tensor(_DevArray(cur_p, (n_cpus, lpc), "<u4"), device=dev).to(torch.int64)
    pos = torch.arange(cap, device=dev)[None, :]
    for c in range(n_cpus):   # one CPU at a time: [lists][cap] weights = 8 MiB
        words = torch.as_tensor(_DevArray(ent_p + c * lpc * cap * 16, (lpc, cap, 4), "<u4"), device=dev)
        w = words[:, :, 3].to(torch.int64)
        out[c] = int((w * (pos < cursors[c][:, None])).sum().item())
    return out


def bind_to_gpu_cpus(local: int):
    """Run this rank (and first-touch its pinned host buffers) on the CPUs next to its GPU:
    with several ranks uploading at once, host buffers on the far socket halve the PCIe rate.
    Returns the CPU list used, or None when NVML has nothing to say."""
    try:
        import pynvml
        pynvml.nvmlInit()
        try:  # CUDA ordinal -> NVML handle through the PCI address (CUDA_VISIBLE_DEVICES may reorder)
            pr = torch.cuda.get_device_properties(local)
            h = pynvml.nvmlDeviceGetHandleByPciBusId(
                f"{pr.pci_domain_id:08x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0")
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByIndex(local)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        near = {64 * i + b for i, w in enumerate(words) for b in range(64) if (int(w) >> b) & 1}
        cpus = sorted(near & os.sched_getaffinity(0))
        if cpus:
            os.sched_setaffinity(0, cpus)
            return cpus
    except Exception:
        pass
    return None


def timed_steps(step, steps, dist, dev):
    """CUDA-event timing of `steps` calls of step(), barrier + sync on both sides, max over ranks."""
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True)
    e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()) / steps


def bench_deltas(args, rank, world, local, dev, dist):
    """Diagnostic workload (config 3 shape): group SAMPLE decode + per-interval deltas."""
    E = args.events
    ring_bytes = args.ring_mib << 20
    mixed = args.workload == "c4_mixed"
    switch = args.workload == "c4_switch"
    hostile = args.workload == "c3_hostile"
    # --shard: args.cpus is the whole machine, sharded by CPU id over the ranks (config 4:
    # "sharded by CPU across 8 B200"); else every rank holds args.cpus rings of its own
    shard = bool(getattr(args, "shard", False)) and world > 1
    cpus = args.cpus // world if shard else args.cpus
    assert cpus > 0
    if switch:
        imgs, tails, heads, n_rec = make_switch_rings(cpus, ring_bytes, rank * cpus, dev)
    elif mixed:
        imgs, tails, heads, n_rec = make_mixed_rings(cpus, ring_bytes, rank * cpus, dev, E)
    elif hostile:
        imgs, tails, heads, n_rec = make_hostile_rings(cpus, ring_bytes, rank * cpus, dev, E)
    else:
        imgs, tails, heads, n_rec = make_group_rings(cpus, ring_bytes, rank * cpus, dev, E)
    window_bytes = sum(h - t for t, h in zip(tails, heads))
    ctx = hbt.Context(local, max_cpus=cpus)
    ctx.use_torch_stream()
    ctx.ring_configure(cpus, 0)
    if args.segs_per_cpu:
        ctx.set_geometry(args.segs_per_cpu)
    if args.kernel:
        ctx.set_decode_kernel({"tiles": hbt.KERNEL_TILES, "walk": hbt.KERNEL_WALK}[args.kernel])
    if dist is not None:
        ctx.comm_init(dist)
    desc = hbt.sampling_desc(E, handled=hbt.MODE_MIXED if mixed else hbt.MODE_SAMPLING)
    if switch:
        desc = hbt.switch_desc(tid_level=1)
        ctx.set_cpu_base(rank * cpus)
    if args.agg == "none":  # diagnostic: boundary discovery + per-type counters only
        desc = hbt.count_desc(hbt.MODE_CONTEXT_SWITCH if switch else hbt.MODE_MIXED if mixed
                              else hbt.MODE_SAMPLING)
    bases_a = np.array([imgs[c].data_ptr() for c in range(cpus)], dtype=np.uint64)
    tails_a, heads_a = np.array(tails, dtype=np.uint64), np.array(heads, dtype=np.uint64)
    kms = []

    def step():
        # every step decodes the same trace again: start from the state after enable(),
        # or the second pass would (correctly) report a timestamp regression
        ctx.reset_state(hbt.INVALID_TS)
        ctx.submit_rings_dev(bases_a, ring_bytes, tails_a, heads_a)
        ctx.decode(desc)
        if dist is not None:
            # per-CPU results of the whole machine on every rank: one NCCL all-gather per step
            step.res_all = ctx.fetch_results_all()
            step.res = step.res_all[rank * cpus:(rank + 1) * cpus]
        else:
            step.res = ctx.fetch_results()
        kms.append(ctx.decode_kernel_ms)

    sampler = sampler_for(local) if rank == 0 else None
    for _ in range(max(args.warmup, 3)):
        step()
    res = step.res
    assert int(res["n_records"].sum()) == n_rec
    n_rows = int(res["n_rows"].sum())
    kms.clear()
    l0 = ctx.launches
    wall0 = time.time()
    ms_step = timed_steps(step, args.steps, dist, dev)
    launches = ctx.launches - l0
    clocks = sampler.window(wall0, time.time()) if rank == 0 else None
    if rank != 0:
        ctx.close()
        return None
    peak, peak_src = load_peaks()
    k_ms = sum(kms) / len(kms)
    algo = window_bytes + n_rows * 8 * (2 + E)
    if switch and args.agg != "none":
        # two passes (proto-events, then the prefix sum that places WriteErrorsEnd events): the
        # span is the whole device-timed step; n_rows counts 16-byte events
        algo = window_bytes + n_rows * 16
        ach = algo / (ms_step * 1e-3) / 1e9
        sw_cb = cpu_baseline_switch(args) if world == 1 and not args.no_cpu_baseline else None
        # the Combinator-tree merge of the per-CPU streams, timed once on its own
        m0, m1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        base, n_ev = C.c_void_p(), C.c_uint64()
        ctx._ck(ctx.lib.hbt_switch_events_merged_dev(ctx.h, C.byref(base), C.byref(n_ev)))  # allocates
        # (a local re-decode: the other ranks have left, so no collective in here)
        ctx.reset_state(hbt.INVALID_TS)
        ctx.submit_rings_dev(bases_a, ring_bytes, tails_a, heads_a)
        ctx.decode(desc)
        ctx.fetch_results()
        torch.cuda.synchronize()
        m0.record()
        ctx._ck(ctx.lib.hbt_switch_events_merged_dev(ctx.h, C.byref(base), C.byref(n_ev)))
        m1.record()
        torch.cuda.synchronize()
        merge_ms = m0.elapsed_time(m1)
        return {"metric": METRIC, "value": n_rec * world / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world,
                "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_step,
                "higher_is_better": True, "scaling": "strong" if shard else "weak", "vs_baseline": None, "dtype": "u64",
                "data": "synthetic",
                "config": {"workload": "c4_switch: mode::ContextSwitch records (32-48 B) -> tagstack::Event stream",
                           "cpus_per_gpu": cpus, "ring_mib": args.ring_mib, "records_per_gpu": n_rec,
                           "events_per_gpu": n_rows, "window_bytes_per_gpu": int(window_bytes),
                           "merge_ms": merge_ms, "merge_levels": max(0, (args.cpus - 1).bit_length()),
                           "l2": "inputs exceed L2; no flush needed", "slow_tiles": ctx.slow_tiles,
                           "reruns": ctx.reruns},
                "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                             "traffic": load_traffic("k_walk<SWITCH_EVENTS>"),
                             "kernel": ("k_decode" if args.kernel == "tiles" else "k_walk") + "<SWITCH_EVENTS> + k_sw_scan/emit",
                             "kernel_ms": ms_step, "decode_kernel_ms": k_ms, "algorithmic_bytes": int(algo),
                             "peak_source": peak_src},
                "gpu_launches": int(launches), "clocks": clocks,
                **({"cpu_baseline": sw_cb} if sw_cb else {})}
    if ctx.reruns:
        # the kernel timer brackets the LAST decode launch only; with re-runs the step is several
        # launches and host round trips: the whole device-timed step is the span
        k_ms = ms_step
    ach = algo / (k_ms * 1e-3) / 1e9
    cb = None
    if world == 1 and not args.no_cpu_baseline and not mixed and not switch and args.agg != "none":
        cb = cpu_baseline_deltas(args, E)
    e2e = None
    if world == 1 and not args.no_e2e and not switch and args.agg != "none":
        e2e = e2e_deltas(args, local, dev, imgs, tails, heads, desc, E, n_rec, n_rows, window_bytes)
    line = {"metric": METRIC, "value": n_rec * world / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "strong" if shard else "weak", "vs_baseline": None, "dtype": "u64",
            "data": "synthetic",
            "config": {"workload": (f"c4_mixed: SAMPLE/SWITCH/COMM/MMAP/LOST/FORK/EXIT stream, per-type counters "
                                    f"+ per-interval deltas ({E} events)") if mixed else
                                   (f"c3_hostile: group SAMPLE ({40 + 8 * E} B) whose counter values are valid-looking record "
                                    "headers (wrong entry guesses are caught and re-run), per-interval deltas") if hostile else
                                   f"c3_deltas: group SAMPLE ({40 + 8 * E} B, {E} events) decode + per-interval deltas",
                       "cpus_per_gpu": cpus, "ring_mib": args.ring_mib, "records_per_gpu": n_rec,
                       "rows_per_gpu": n_rows, "window_bytes_per_gpu": int(window_bytes),
                       "l2": "inputs exceed L2; no flush needed", "slow_tiles": ctx.slow_tiles,
                       "reruns": ctx.reruns, "rerun_segments": ctx.rerun_segments},
            "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                         "traffic": (load_traffic(f"k_decode<COUNT_DELTAS>@{E}ev") if not mixed
                                     else load_traffic("k_walk<COUNT_DELTAS>@mixed")),
                         "kernel": ("k_walk<COUNT_DELTAS>" if (mixed and args.kernel != "tiles") or args.kernel == "walk"
                                    else "k_decode<COUNT_DELTAS>"),
                         "kernel_ms": k_ms, "algorithmic_bytes": int(algo), "peak_source": peak_src},
            "gpu_launches": int(launches), "clocks": clocks}
    if cb:
        line["cpu_baseline"] = cb
    if e2e:
        line["e2e"] = e2e
    return line


def e2e_deltas(args, local, dev, imgs, tails, heads, desc, E, n_rec, n_rows, window_bytes):
    """End to end for the deltas path: ring images in pinned HOST memory through hbt_submit_ring,
    decode, per-CPU results and ALL delta rows back to pinned host memory (hbt_fetch_rows_all: the
    slabs are packed on the device and leave in one copy per context).  CPU groups are pipelined
    over PCIe like the headline's e2e leg: one hbt_ctx + stream per group."""
    cpus, ring_bytes = len(tails), args.ring_mib << 20
    host = torch.empty((cpus, ring_bytes // 8), dtype=torch.int64).pin_memory()
    host.copy_(imgs)
    host_np = host.numpy()
    G = max(1, min(args.e2e_groups, cpus))
    per = (cpus + G - 1) // G
    groups = []
    for g in range(G):
        lo, hi = g * per, min(cpus, (g + 1) * per)
        if lo >= hi:
            break
        cg = hbt.Context(local, max_cpus=hi - lo)
        sg = torch.cuda.Stream(dev)
        cg.set_stream(sg.cuda_stream)
        cg.ring_configure(hi - lo, ring_bytes)
        cap_rows = int(n_rows * 1.1 * (hi - lo) / cpus) + 1024
        pinned = torch.empty((cap_rows, 2 + E), dtype=torch.int64).pin_memory()
        groups.append((cg, lo, hi, pinned.numpy().view(np.uint64), sg, pinned))

    def step():
        uploaded = None
        for cg, lo, hi, _, sg, _ in groups:
            cg.reset_state(hbt.INVALID_TS)
            if uploaded is not None:
                sg.wait_event(uploaded)
            for c in range(lo, hi):
                cg.submit_ring_image(c - lo, host_np[c], tails[c], heads[c])
            uploaded = torch.cuda.Event()
            uploaded.record(sg)
            cg.decode(desc)
        recs = rows = 0
        for cg, lo, hi, buf, _, _ in groups:
            recs += int(cg.fetch_results()["n_records"].sum())
            rows += len(cg.fetch_rows_all(E, out=buf)[0])
        return recs, rows

    step()
    torch.cuda.synchronize()
    steps = max(1, min(args.steps, 5))
    t0 = time.perf_counter()
    for _ in range(steps):
        recs, rows = step()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / steps
    assert recs == n_rec and rows == n_rows, (recs, n_rec, rows, n_rows)
    for cg, *_ in groups:
        cg.close()
    return {"value": n_rec / dt, "unit": UNIT, "h2d_bytes_per_step": int(window_bytes),
            "d2h_bytes_per_step": int(n_rows * 8 * (2 + E)), "ms_per_step": dt * 1e3, "steps": steps,
            "pipeline": f"{len(groups)} CPU groups, one hbt_ctx + stream each; rows packed on the device, "
                        "one device-to-host copy per group"}


def bench_counts(args, rank, world, local, dev, dist):
    """Config 5 shape: counter-group reduce over CPUs, scaling + rates + IPC ratio + interval
    deltas.  At N > 1 the args.cpus CPUs are SHARDED over the ranks (strong scaling over CPUs, the
    north star's "512-CPU trace ... merged via NCCL reduce"): every timed step is
    hbt_counts_pipeline_sharded_dev = local reduce -> ncclAllReduce(u64 sum) of the [I][3+E]
    partial sums -> scale, in interval chunks with the collective on a second stream."""
    from dynolog_1_b200 import sharding
    E, C_total, I = args.events, args.cpus, args.intervals
    mine = sharding.shard_cpus(C_total, world, rank)
    C = len(mine)
    assert C > 0, "more ranks than CPUs"
    # (seeded per CPU block so that a rank's shard does not depend on the world size)
    reads = tg.counts_trace(tg.stream_seed(5, rank), I, C, E, device=dev, multiplexed=True)
    ctx = hbt.Context(local, max_cpus=max(C, 1))
    ctx.use_torch_stream()
    if dist is not None:
        ctx.comm_init(dist)
    W = 3 + E
    sums = torch.zeros((I, W), dtype=torch.int64, device=dev)
    scaled = torch.zeros((I, E), dtype=torch.int64, device=dev)
    rates = torch.zeros((I, E), dtype=torch.float32, device=dev)
    ratio = torch.zeros((I, 1), dtype=torch.float64, device=dev)
    deltas = torch.zeros((I, E), dtype=torch.int64, device=dev)
    ratios = [(min(1, E - 1), 0)]

    def step():
        if dist is None:
            ctx.counts_pipeline_dev(reads, I, C, E, ratios, sums, scaled, rates, ratio, deltas)
        else:
            ctx.counts_pipeline_sharded_dev(reads, I, C, E, ratios, sums, scaled, rates, ratio, deltas)

    def step_local():  # the same work without the exchange: what the collective costs on top
        ctx.counts_reduce_dev(reads, I, C, E, sums)
        ctx.counts_scale_dev(sums, I, E, ratios, None, scaled, rates, ratio, deltas)

    sampler = sampler_for(local) if rank == 0 else None
    for _ in range(max(args.warmup, 3)):
        step()
    ctx.sync()
    l0 = ctx.launches
    wall0 = time.time()
    ms_step = timed_steps(step, args.steps, dist, dev)
    launches = ctx.launches - l0
    clocks = sampler.window(wall0, time.time()) if rank == 0 else None
    ms_local = None
    # cross-check (outside the timed region): the merged sums' time_enabled column is the sum
    # over ALL CPUs -- every rank must hold the same table
    if dist is not None:
        chk = sums[:, 1].sum().reshape(1).clone()
        lo, hi = chk.clone(), chk.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        assert int(lo.item()) == int(hi.item()), "ranks disagree on the merged sums"
    if dist is not None:
        step_local()
        ms_local = timed_steps(step_local, args.steps, dist, dev)
    ctx.close()
    if rank != 0:
        return None
    peak, peak_src = load_peaks()
    # algorithmic bytes of the whole job: every rank reads its shard once, the outputs are written once
    algo = I * C_total * W * 8 + I * (W * 8 + E * 8 + E * 4 + 8 + E * 8)
    ach = algo / world / (ms_step * 1e-3) / 1e9   # per GPU: the roofline is a per-GPU peak
    line = {"metric": "counter-group reads decoded+aggregated/sec", "value": I * C_total / (ms_step * 1e-3),
            "unit": "reads/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong" if world > 1 else "weak",
            "vs_baseline": None, "dtype": "u64/f64", "data": "synthetic",
            "config": {"workload": f"c5_counts: {C_total} CPUs x {E} events x {I} intervals"
                                   + (f", sharded {C} CPUs per GPU: reduce -> ncclAllReduce(u64) -> scale/rates/ratio/deltas"
                                      if world > 1 else ": reduce -> scale/rates/ratio/deltas"),
                       "cpus_per_gpu": C,
                       "l2": f"input {I * C * W * 8 / 2**20:.0f} MiB per GPU per step exceeds L2" if I * C * W * 8 > 2**27
                             else "input smaller than L2: not a valid HBM measurement"},
            "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                         "traffic": load_traffic("k_counts_reduce"), "kernel": "k_counts_reduce (+ scale, diff)",
                         "kernel_ms": ms_step, "algorithmic_bytes": int(algo // world), "peak_source": peak_src,
                         "note": "per GPU, timed as the whole step (reduce, [all-reduce], scale, diff), not one kernel"},
            "gpu_launches": int(launches), "clocks": clocks}
    if ms_local is not None:
        line["collective"] = {"op": "ncclAllReduce(u64, sum) of [I][3+E] partial sums, chunked, on a second stream",
                              "bytes_per_step": int(I * W * 8), "ms_step_without_exchange": ms_local,
                              "share_of_step": max(0.0, 1.0 - ms_local / ms_step)}
    return line


def run_reference(args, rank, world):
    """The reference's consume() (reference headers, oracle/_ref) + the harness's
    (pid,ip) map, all host threads, on a bounded sample of the same workload."""
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _oracle as orc
    import ctypes as C
    cores = os.cpu_count() or 1
    kind = "reference"
    if not orc.have_ref():
        print(json.dumps({"impl": "reference", "unavailable":
                          "oracle/_ref/libhbt_ref.so was not built (no reference tree at build time)"}))
        return
    lib = orc.ref()
    ring_bytes = args.ring_mib << 20
    # bounded sample: `sample_cpus` rings of the 256 (same generator, same seeds)
    sample_cpus = min(args.cpus, args.ref_cpus)
    imgs, tails, heads, n_rec = make_c2(sample_cpus, ring_bytes, 0, "cpu")
    rings = imgs.numpy().view(np.uint8)
    tails_a = np.array(tails, dtype=np.uint64)
    heads_a = np.array(heads, dtype=np.uint64)
    res = (hbt.CpuResult * sample_cpus)()
    h = lib.hbtref_hist_new()
    times = []
    for it in range(args.warmup + args.steps):
        for r in res:
            for i in range(hbt.NUM_COUNTERS):
                r.counters[i] = 0
        t = lib.hbtref_ip_hist(h, rings.ctypes.data, ring_bytes, tails_a.ctypes.data,
                               heads_a.ctypes.data, sample_cpus, 40, hbt.DRAIN, cores, res)
        assert t > 0, t
        if it >= args.warmup:
            times.append(t)
    got = sum(int(r.counters[hbt.CNT_SAMPLE]) for r in res)
    assert got == n_rec, (got, n_rec)
    lib.hbtref_hist_free(h)
    dt = sum(times) / len(times)
    value = n_rec / dt
    sample = f"{sample_cpus} of {args.cpus} rings x {args.ring_mib} MiB per step"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64",
            "data": "synthetic",
            "config": {"workload": "c2_ip_hist: SAMPLE ring decode + (cpu,pid,ip) histogram",
                       "cpus": args.cpus, "ring_mib": args.ring_mib, "sample_bytes": 40,
                       "note": "CPU reference arm: bounded sample of the workload per step"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
                             "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def cpu_baseline_deltas(args, n_events: int):
    """The reference's own CpuCountSampleGenerator (oracle/_ref) over a bounded sample of the
    c3_deltas rings, one generator per ring, all host cores."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    try:
        import _oracle as orc
        if not orc.have_ref():
            return None
        lib = orc.ref()
    except Exception:
        return None
    cores = os.cpu_count() or 1
    ring_bytes = args.ring_mib << 20
    sample_cpus = min(args.cpus, args.ref_cpus)
    imgs, tails, heads, n_rec = make_group_rings(sample_cpus, ring_bytes, 0, "cpu", n_events)
    rings = imgs.numpy().view(np.uint8)
    tails_a, heads_a = np.array(tails, dtype=np.uint64), np.array(heads, dtype=np.uint64)
    lib.hbtref_sample_deltas_timed.restype = C.c_double
    lib.hbtref_sample_deltas_timed.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint32,
                                               C.c_uint32, C.c_uint32, C.POINTER(C.c_uint64),
                                               C.POINTER(C.c_uint64)]
    best = None
    recs, rows = C.c_uint64(), C.c_uint64()
    # the reference logs a warning per LOST record (PerCpuCountSampleGenerator.h:111): keep it
    # (it is part of what the reference does) but not on this process's stderr
    sys.stderr.flush()
    saved, devnull = os.dup(2), os.open(os.devnull, os.O_WRONLY)
    os.dup2(devnull, 2)
    try:
        for _ in range(2):
            t = lib.hbtref_sample_deltas_timed(rings.ctypes.data, ring_bytes, tails_a.ctypes.data,
                                               heads_a.ctypes.data, sample_cpus, n_events, cores,
                                               C.byref(recs), C.byref(rows))
            best = t if best is None else min(best, t)
    finally:
        os.dup2(saved, 2)
        os.close(saved)
        os.close(devnull)
    return {"value": n_rec / best, "unit": UNIT, "cores": cores, "kind": "reference",
            "sample": f"{sample_cpus} of {args.cpus} rings x {args.ring_mib} MiB, best of 2; the reference's "
                      "CpuCountSampleGenerator::consume() with its own output ring",
            "rows": int(rows.value)}


def cpu_baseline_switch(args):
    """The reference's own CpuThreadSwitchGenerator (oracle/_ref) over a bounded sample of the
    c4_switch rings.  One thread: its generators share one unsynchronised PerThreadStats map
    (thread ids -> virtual ids, with a /proc lookup per new thread), as in the reference."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    try:
        import _oracle as orc
        if not orc.have_ref():
            return None
    except Exception:
        return None
    ring_bytes = args.ring_mib << 20
    sample_cpus = min(args.cpus, 2)  # ~4 us per record (a /proc lookup per new thread): keep it short
    imgs, tails, heads, n_rec = make_switch_rings(sample_cpus, ring_bytes, 0, "cpu")
    rings = imgs.numpy().view(np.uint64)
    best = None
    for _ in range(2):
        t0 = time.perf_counter()
        ref = orc.ref_switch_events(rings, tails, heads, tid_level=1)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return {"value": n_rec / best, "unit": UNIT, "cores": 1, "kind": "reference",
            "sample": f"{sample_cpus} of {args.cpus} rings x {args.ring_mib} MiB, best of 2; the reference's "
                      "CpuThreadSwitchGenerator::consume() with its own output ring and thread map",
            "events": int(sum(len(r["events"]) for r in ref))}


def cpu_baseline_leg(args, gpu_sample=None):
    """Bounded CPU baseline beside the GPU number (rank 0, N=1 only): the reference's consume()
    (oracle/_ref) + the harness's (pid,ip) map over the first ref_cpus rings.  gpu_sample =
    (GPU histogram entries of those CPUs, per-CPU GPU results): when given, the reference's
    histogram and per-CPU counters of the sample are compared with them entry by entry and the
    outcome is reported as "parity_check"."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    try:
        import _oracle as orc
        if not orc.have_ref():
            return None
        lib = orc.ref()
    except Exception:
        return None
    cores = os.cpu_count() or 1
    ring_bytes = args.ring_mib << 20
    sample_cpus = min(args.cpus, args.ref_cpus)
    imgs, tails, heads, n_rec = make_c2(sample_cpus, ring_bytes, 0, "cpu")
    rings = imgs.numpy().view(np.uint8)
    tails_a = np.array(tails, dtype=np.uint64)
    heads_a = np.array(heads, dtype=np.uint64)
    res = (hbt.CpuResult * sample_cpus)()
    h = lib.hbtref_hist_new()
    best = None
    for _ in range(2):
        for r in res:
            for i in range(hbt.NUM_COUNTERS):
                r.counters[i] = 0
        t = lib.hbtref_ip_hist(h, rings.ctypes.data, ring_bytes, tails_a.ctypes.data,
                               heads_a.ctypes.data, sample_cpus, 40, hbt.DRAIN, cores, res)
        best = t if best is None else min(best, t)
    parity = None
    if gpu_sample is not None:
        got_ent, got_res = gpu_sample
        m = int(lib.hbtref_hist_size(h))
        ref_ent = np.zeros(max(1, m), dtype=hbt.HIST_DTYPE)
        k = int(lib.hbtref_hist_fetch(h, ref_ent.ctypes.data, ref_ent.shape[0]))
        ref_ent = ref_ent[:k]

        def canon(e):
            o = np.lexsort((e["ip"], e["pid"], e["cpu"]))
            return e[o]
        a, b = canon(got_ent), canon(ref_ent)
        same_hist = len(a) == len(b) and bool((a == b).all())
        same_res = all(int(got_res[c]["ret"]) == int(res[c].ret) and int(got_res[c]["n_records"]) == int(res[c].n_records)
                       and int(got_res[c]["new_tail"]) == int(res[c].new_tail)
                       and int(res[c].counters[hbt.CNT_SAMPLE]) == int(got_res[c]["n_records"])
                       for c in range(sample_cpus))
        parity = {"result": "ok" if (same_hist and same_res) else "MISMATCH",
                  "checked": f"(cpu,pid,ip) histogram of CPUs 0-{sample_cpus - 1} ({len(b)} keys, "
                             f"{int(b['count'].sum())} samples) and their consume() results, GPU vs reference, bit-exact"}
    lib.hbtref_hist_free(h)
    # one thread, as the reference runs it (one collecting thread), on 4 of the rings
    one = min(4, sample_cpus)
    h1 = lib.hbtref_hist_new()
    t1 = lib.hbtref_ip_hist(h1, rings.ctypes.data, ring_bytes, tails_a.ctypes.data, heads_a.ctypes.data,
                            one, 40, hbt.DRAIN, 1, res)
    lib.hbtref_hist_free(h1)
    out = {"value": n_rec / best, "unit": UNIT, "cores": cores, "kind": "reference",
           "sample": f"{sample_cpus} of {args.cpus} rings x {args.ring_mib} MiB, best of 2; "
                     "reference consume() + std::unordered_map handler",
           "single_thread_value": (n_rec // sample_cpus) * one / t1 if t1 > 0 else None}
    return out, parity


def c1_reference_leg(args):
    """BASELINE configs[0], the reference's own CPU-runnable case: PerCpuCountReader-style group
    reads + getCount + PerfMonitor's rates over a 64-CPU x 8-event x 10k-interval trace, through
    the reference's PerCpuBase::read / GroupReadValues (oracle/_ref), all host cores."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    try:
        import _oracle as orc
        if not orc.have_ref():
            return None
    except Exception:
        return None
    I, Cc, E = 10000, 64, 8
    reads = tg.counts_trace(tg.stream_seed(1, 0), I, Cc, E, multiplexed=False).numpy().view(np.uint64)
    cores = os.cpu_count() or 1
    best = None
    for _ in range(3):
        o = orc.ref_counts(reads, [(1, 0)], threads=cores)
        best = o["seconds"] if best is None else min(best, o["seconds"])
    one = orc.ref_counts(reads, [(1, 0)], threads=1)["seconds"]
    return {"workload": "c1: 64 CPUs x 8 events x 10k intervals, PerCpuBase::read + getCount + rates + IPC (CPU reference, no GPU)",
            "value": I * Cc / best, "unit": "reads/s", "cores": cores, "kind": "reference", "seconds": best,
            "single_thread_value": I * Cc / one, "sample": "the whole config, best of 3"}


def run_diagnostics(args, rank, world, local, dev, dist):
    """The other BASELINE configs beside the headline, a few steps each, every one with its own
    clock sample.  N = 1: C3 deltas (32 events), C4 mixed stream, C4 switch events, C5 counter
    groups (512 CPUs x 1 event x 1M intervals, and the 64 x 8 shape), C1 on the host cores.
    N > 1: the configs the north star shards -- C5 over 512 CPUs and C4's mixed stream over 256
    CPUs, split over the ranks, with the NCCL exchange inside every timed step."""
    import copy
    import gc

    def variant(**kw):
        a = copy.copy(args)
        for k, v in kw.items():
            setattr(a, k, v)
        return a

    if world == 1:
        specs = [("c3_deltas_32ev", bench_deltas, variant(workload="c3_deltas", events=32, agg="hist")),
                 ("c3_deltas_8ev", bench_deltas, variant(workload="c3_deltas", events=8, agg="hist", no_cpu_baseline=True, no_e2e=True)),
                 ("c4_mixed", bench_deltas, variant(workload="c4_mixed", events=8, agg="hist", no_e2e=True)),
                 ("c4_switch", bench_deltas, variant(workload="c4_switch", agg="hist")),
                 ("c3_hostile_payload", bench_deltas, variant(workload="c3_hostile", events=8, agg="hist",
                                                              no_cpu_baseline=True, no_e2e=True)),
                 ("c2_count_only", bench_c2_variant, variant(c2_agg="none", c2_words=5)),
                 ("c2_ip_hist_128B_samples", bench_c2_variant, variant(c2_agg="hist", c2_words=16)),
                 ("c5_counts_512x1", bench_counts, variant(workload="c5_counts", cpus=512, events=1, intervals=1_000_000)),
                 ("c5_counts_64x8", bench_counts, variant(workload="c5_counts", cpus=64, events=8, intervals=400_000))]
    else:
        specs = [("c5_counts_512x1_sharded", bench_counts,
                  variant(workload="c5_counts", cpus=512, events=1, intervals=1_000_000)),
                 ("c4_mixed_256cpus_sharded", bench_deltas,
                  variant(workload="c4_mixed", events=8, cpus=256, shard=True, agg="hist", no_cpu_baseline=True))]
    out = {}
    for name, fn, a in specs:
        a.steps = max(args.diag_steps, 3)
        gc.collect()
        torch.cuda.empty_cache()
        try:
            line = fn(a, rank, world, local, dev, dist)
        except Exception as e:  # a diagnostic must not take the headline down with it
            line = {"error": f"{type(e).__name__}: {e}"}
            if dist is not None:
                raise
        if rank != 0 or line is None:
            continue
        keep = {k: line[k] for k in ("value", "unit", "ms_per_step", "steps", "scaling", "clocks", "cpu_baseline",
                                     "collective", "gpu_launches", "error", "e2e") if k in line}
        if "config" in line:
            keep["workload"] = line["config"].get("workload")
            for k in ("cpus_per_gpu", "slow_tiles", "reruns", "rerun_segments", "merge_ms", "l2"):
                if k in line["config"]:
                    keep[k] = line["config"][k]
        if "roofline" in line:
            keep["roofline"] = {k: line["roofline"][k] for k in ("achieved", "peak", "frac", "kernel", "kernel_ms",
                                                                 "algorithmic_bytes") if k in line["roofline"]}
        out[name] = keep
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        c1 = c1_reference_leg(args)
        if c1:
            out["c1_cpu_reference"] = c1
    return out


def bench_c2_variant(args, rank, world, local, dev, dist):
    """The headline's rings, device-resident only, in a variant: args.c2_agg = "none" decodes
    with per-type counters only (how fast the boundary discovery alone streams the 4 GiB);
    args.c2_words = 16 uses 128-byte samples (sample_type IP|TID|TIME|CPU|READ with an 8-event
    read group: SURVEY 8d allows "40 B + optional read group") with the histogram."""
    ring_bytes = args.ring_mib << 20
    words = getattr(args, "c2_words", 5)
    hist = getattr(args, "c2_agg", "hist") == "hist"
    imgs, tails, heads, n_rec = make_c2(args.cpus, ring_bytes, rank * args.cpus, dev, sample_words=words)
    window_bytes = sum(h - t for t, h in zip(tails, heads))
    ctx = hbt.Context(local, max_cpus=args.cpus)
    ctx.use_torch_stream()
    ctx.ring_configure(args.cpus, 0)
    if hist:
        ctx.hist_configure(args.hist_slots_per_cpu * args.cpus)
    desc = hbt.ip_desc(8 * words, aggregate=hbt.AGG_IP_HISTOGRAM if hist else hbt.AGG_NONE)
    bases_a = np.array([imgs[c].data_ptr() for c in range(args.cpus)], dtype=np.uint64)
    tails_a, heads_a = np.array(tails, dtype=np.uint64), np.array(heads, dtype=np.uint64)
    kms = []

    def step():
        if hist:
            ctx.hist_clear()
        ctx.submit_rings_dev(bases_a, ring_bytes, tails_a, heads_a)
        ctx.decode(desc)
        step.res = ctx.fetch_results()
        kms.append(ctx.decode_kernel_ms + (ctx.hist_fold_kernel_ms if hist else 0.0))

    sampler = sampler_for(local) if rank == 0 else None
    for _ in range(3):
        step()
    assert int(step.res["n_records"].sum()) == n_rec
    distinct = ctx.hist_size() if hist else 0
    kms.clear()
    wall0 = time.time()
    ms_step = timed_steps(step, args.steps, dist, dev)
    clocks = sampler.window(wall0, time.time()) if rank == 0 else None
    ctx.close()
    if rank != 0:
        return None
    peak, _ = load_peaks()
    k_ms = sum(kms) / len(kms)
    algo = window_bytes + 16 * distinct
    ach = algo / (k_ms * 1e-3) / 1e9
    what = (f"c2 rings with {8 * words}-byte samples, decode + (cpu,pid,ip) histogram" if hist
            else "c2 rings, decode + per-type counters only (HBT_AGG_NONE)")
    return {"value": n_rec * world / (ms_step * 1e-3), "unit": UNIT, "ms_per_step": ms_step, "steps": args.steps,
            "scaling": "weak", "clocks": clocks,
            "config": {"workload": what, "cpus_per_gpu": args.cpus},
            "roofline": {"achieved": ach, "peak": peak, "frac": ach / peak,
                         "kernel": "k_decode<IP_HISTOGRAM> + k_hist_aggregate" if hist else "k_decode<NONE>",
                         "kernel_ms": k_ms, "algorithmic_bytes": int(algo)}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cpus", type=int, default=256)
    ap.add_argument("--ring-mib", type=int, default=16)
    ap.add_argument("--ref-cpus", type=int, default=32, help="rings in the bounded CPU sample")
    ap.add_argument("--hist-slots-per-cpu", type=int, default=1 << 21,
                    help="list capacity per CPU: distinct keys + the samples of one decode")
    ap.add_argument("--agg", default="hist", choices=["hist", "none"],
                    help="none: decode + counters only (diagnostic, not the headline workload)")
    ap.add_argument("--workload", default="c2_ip_hist", choices=["c2_ip_hist", "c3_deltas", "c3_hostile", "c4_mixed", "c4_switch", "c5_counts"],
                    help="c2_ip_hist is the headline (BASELINE configs[1]); the others are diagnostics")
    ap.add_argument("--events", type=int, default=8)
    ap.add_argument("--intervals", type=int, default=100000)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-groups", type=int, default=8,
                    help="CPU groups (contexts) the end-to-end leg pipelines over PCIe")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-diagnostics", action="store_true",
                    help="skip the diagnostics block (the other BASELINE configs)")
    ap.add_argument("--diag-steps", type=int, default=20, help="timed steps of each diagnostic workload")
    ap.add_argument("--segs-per-cpu", type=int, default=0, help="hbt_set_geometry (0 = automatic)")
    ap.add_argument("--kernel", default="", choices=["", "tiles", "walk"], help="hbt_set_decode_kernel (default: automatic)")
    ap.add_argument("--shard", action="store_true",
                    help="ring workloads at N > 1: --cpus is the whole machine, sharded over the ranks")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    assert torch.cuda.is_available(), "bench.py needs a GPU (there is no CPU fallback)"
    torch.cuda.set_device(local)
    all_cpus = os.sched_getaffinity(0)
    numa = bind_to_gpu_cpus(local)
    dev = torch.device("cuda", local)
    # everything (torch ops, the library's kernels, the timing events) runs on ONE
    # non-default stream, so torch.cuda.Event sees the kernels it brackets
    torch.cuda.set_stream(torch.cuda.Stream(dev))
    dist = None
    if world > 1:
        import torch.distributed as dist_
        dist = dist_
        dist.init_process_group("nccl", device_id=dev)

    if args.workload != "c2_ip_hist":
        fn = bench_counts if args.workload == "c5_counts" else bench_deltas
        line = fn(args, rank, world, local, dev, dist)
        if rank == 0:
            print(json.dumps(line))
        if dist is not None:
            dist.destroy_process_group()
        return

    sampler = sampler_for(local) if rank == 0 else None  # started long before the timed region
    ring_bytes = args.ring_mib << 20
    cpu_base = rank * args.cpus
    imgs, tails, heads, n_rec = make_c2(args.cpus, ring_bytes, cpu_base, dev)
    window_bytes = sum(h - t for t, h in zip(tails, heads))
    torch.cuda.synchronize()

    ctx = hbt.Context(local, max_cpus=args.cpus)
    ctx.use_torch_stream()
    ctx.ring_configure(args.cpus, 0)
    ctx.set_cpu_base(cpu_base)
    ctx.hist_configure(args.hist_slots_per_cpu * args.cpus)
    if args.segs_per_cpu:
        ctx.set_geometry(args.segs_per_cpu)
    if args.kernel:
        ctx.set_decode_kernel({"tiles": hbt.KERNEL_TILES, "walk": hbt.KERNEL_WALK}[args.kernel])
    if dist is not None:
        ctx.comm_init(dist)   # NCCL communicator of the library itself (hbt_comm_init)
    desc = hbt.ip_desc(40, aggregate=hbt.AGG_IP_HISTOGRAM if args.agg == "hist" else hbt.AGG_NONE)

    bases_a = np.array([imgs[c].data_ptr() for c in range(args.cpus)], dtype=np.uint64)
    tails_a = np.array(tails, dtype=np.uint64)
    heads_a = np.array(heads, dtype=np.uint64)

    def step():
        if args.agg == "hist":
            ctx.hist_clear()
        ctx.submit_rings_dev(bases_a, ring_bytes, tails_a, heads_a)
        ctx.decode(desc)

    # ---- warm-up + validation -------------------------------------------------
    for it in range(max(args.warmup, 3)):
        step()
        if dist is not None:   # the same calls as the timed steps (the first gather sets NCCL up)
            res = ctx.fetch_results_all()[rank * args.cpus:(rank + 1) * args.cpus]
        else:
            res = ctx.fetch_results()
        # counters are running totals since reset_state, like the reference's members
        assert int(res["counters"][:, hbt.CNT_SAMPLE].sum()) == n_rec * (it + 1)
    total = int(res["n_records"].sum())
    assert total == n_rec, (total, n_rec)
    distinct = ctx.hist_size() if args.agg == "hist" else 0
    slow, reruns = ctx.slow_tiles, ctx.reruns
    gpu_sample = None
    if args.agg == "hist":
        # every sample must be in the table exactly once: weights summed over each CPU's lists on
        # the device == that CPU's record count (all records are samples)
        per_cpu = hist_weight_per_cpu(ctx, args.cpus, dev)
        assert np.array_equal(per_cpu, res["n_records"].astype(np.int64)), "histogram weights != samples per CPU"
        if rank == 0 and world == 1 and not args.no_cpu_baseline:
            # kept for the parity check against the reference's histogram of the same CPUs
            gpu_sample = (ctx.hist_fetch_cpus(0, min(args.cpus, args.ref_cpus)), res.copy())

    # ---- timed region: device-resident ----------------------------------------
    launches0 = ctx.launches
    kernel_ms, fold_ms = [], []
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True)
    e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    wall0 = time.time()
    host_us = []
    for _ in range(args.steps):
        h0 = time.perf_counter()
        step()
        host_us.append((time.perf_counter() - h0) * 1e6)  # host time to queue one step
        # completing the decode (re-run check) is part of the step; it syncs the stream.  The
        # decode itself shards by CPU with no exchange; at N > 1 the step ends with every rank's
        # per-CPU results gathered on every rank (one NCCL all-gather, hbt_fetch_results_all)
        last_res = ctx.fetch_results_all() if dist is not None else ctx.fetch_results()
        kernel_ms.append(ctx.decode_kernel_ms)
        if args.agg == "hist":
            fold_ms.append(ctx.hist_fold_kernel_ms)
    e1.record()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    ms_total = e0.elapsed_time(e1)
    launches = ctx.launches - launches0
    clocks = sampler.window(wall0, time.time()) if rank == 0 else None
    # the gathered table holds every rank's CPUs: running SAMPLE counters of the whole machine
    merged_samples = int(last_res["counters"][:, hbt.CNT_SAMPLE].sum())
    assert len(last_res) == args.cpus * world
    assert merged_samples == n_rec * world * (max(args.warmup, 3) + args.steps), merged_samples

    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    # per-rank spread (diagnostic): slowest rank's kernel span and host queueing time per step
    spread = torch.tensor([(sum(kernel_ms) + sum(fold_ms)) / len(kernel_ms), sum(host_us) / len(host_us)],
                          dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(spread, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / args.steps
    value = n_rec * world / (ms_step * 1e-3)

    # ---- end-to-end: host buffers through the C-ABI --------------------------------
    e2e = None
    if not args.no_e2e:
        host = torch.empty((args.cpus, ring_bytes // 8), dtype=torch.int64).pin_memory()
        host.copy_(imgs)
        host_np = host.numpy()
        ctx.close()
        del imgs
        torch.cuda.empty_cache()
        # The CPUs are handled in G groups, one hbt_ctx (own stream) each — the reference's
        # "one generator set per collecting thread" — so the H2D of group g+1 overlaps the
        # decode, fold and D2H of group g (PCIe is full duplex; the copies dominate).
        G = max(1, min(args.e2e_groups, args.cpus))
        per = (args.cpus + G - 1) // G
        # group boundaries: G groups of `per` CPUs, the last one tapered (halved down to 4 CPUs) -- what
        # cannot overlap anything is the decode + download of the LAST group, so it is kept small
        bounds, lo = [], 0
        while lo < args.cpus:
            hi = min(args.cpus, lo + per)
            if hi == args.cpus:
                while hi - lo > 4:
                    mid = lo + (hi - lo + 1) // 2
                    bounds.append((lo, mid))
                    lo = mid
            bounds.append((lo, hi))
            lo = hi
        groups = []
        ent_dt = hbt.HIST_PACKED_DTYPE.itemsize   # 16-byte entries + a per-CPU offset table (hbt_hist_fetch_packed)
        for lo, hi in bounds:
            cg = hbt.Context(local, max_cpus=hi - lo)
            sg = torch.cuda.Stream(dev)
            cg.set_stream(sg.cuda_stream)
            cg.ring_configure(hi - lo, ring_bytes)
            cg.set_cpu_base(cpu_base + lo)
            cg.hist_configure(args.hist_slots_per_cpu * (hi - lo))
            # result buffer in pinned host memory (a pageable destination makes the D2H
            # copy of ~1 GB of entries the slowest part of the step)
            cap = int(distinct * 1.25 * (hi - lo) / args.cpus) + 4096
            pinned = torch.empty(cap * ent_dt, dtype=torch.uint8).pin_memory()
            groups.append((cg, lo, hi, pinned.numpy().view(hbt.HIST_PACKED_DTYPE), (pinned, sg)))

        def e2e_step():
            uploaded = None
            for cg, lo, hi, _, (_, sg) in groups:
                cg.hist_clear()
                cg.reset_state(hbt.INVALID_TS)
                # uploads of different streams would otherwise share the link round-robin and
                # all finish together: chain them so group g is complete before g+1 starts
                # (decode() itself issues no copy-engine work, so nothing of group g queues
                # behind the next group's upload)
                if uploaded is not None:
                    sg.wait_event(uploaded)
                for c in range(lo, hi):
                    cg.submit_ring_image(c - lo, host_np[c], tails[c], heads[c])
                uploaded = torch.cuda.Event()
                uploaded.record(sg)
                cg.decode(desc)
            rs, n_ent = [], 0
            trace = [time.perf_counter()]
            for cg, lo, hi, buf, _ in groups:
                rs.append(cg.fetch_results())
                trace.append(time.perf_counter())
                n_ent += len(cg.hist_fetch_packed(0, hi - lo, out=buf)[0])
                trace.append(time.perf_counter())
            if os.environ.get("HBT_BENCH_TRACE"):
                print("e2e trace (ms since issue):",
                      " ".join(f"{(x - trace[0]) * 1e3:.1f}" for x in trace[1:]), file=sys.stderr)
            return np.concatenate(rs), n_ent

        e2e_step()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        e2e_steps = max(1, args.steps)   # the requested number of steps, like the device-resident leg
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            r, n_ent = e2e_step()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / e2e_steps
        tt = torch.tensor([dt], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt.item())
        assert int(r["n_records"].sum()) == n_rec
        assert n_ent == distinct, (n_ent, distinct)

        # The floor of that step on this host: the same uploads with nothing else (no decode, no
        # download), all ranks at once -- what the PCIe links and the host's memory system give
        # when every GPU of the box pulls its rings at the same time.
        def upload_only():
            for cg, lo, hi, _, (_, sg) in groups:
                for c in range(lo, hi):
                    cg.submit_ring_image(c - lo, host_np[c], tails[c], heads[c])
            for cg, *_ in groups:
                cg.sync()
        upload_only()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(3):
            upload_only()
        up = (time.perf_counter() - t0) / 3
        tu = torch.tensor([up], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(tu, op=dist.ReduceOp.MAX)
        up = float(tu.item())
        e2e = {"value": n_rec * world / dt, "unit": UNIT, "h2d_bytes_per_step": int(window_bytes),
               "d2h_bytes_per_step": int(n_ent * ent_dt + 8 * (args.cpus + len(groups)) + r.nbytes), "ms_per_step": dt * 1e3,
               "steps": e2e_steps,
               "upload_only_ms_per_step": up * 1e3,
               "upload_only_gbs_per_gpu": window_bytes / up / 1e9,
               "frac_of_upload_floor": up / dt,
               "pipeline": f"{len(groups)} CPU groups ({per} CPUs each, the last one tapered down to 4), one hbt_ctx + "
                           "stream each: H2D of group g+1 overlaps decode + fold + D2H of group g",
               "host_cpus_bound": len(numa) if numa else None}
        for cg, *_ in groups:
            cg.close()
        del host, host_np, groups

    # ---- the other BASELINE configs (all ranks take part: the sharded ones hold collectives) ----
    diagnostics = None
    if not args.no_diagnostics:
        try:
            ctx.close()
        except Exception:
            pass
        imgs = None
        diagnostics = run_diagnostics(args, rank, world, local, dev, dist)

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    peak, peak_src = load_peaks()
    k_ms = sum(kernel_ms) / len(kernel_ms)
    f_ms = sum(fold_ms) / len(fold_ms) if fold_ms else 0.0
    span_ms = k_ms + f_ms  # decode + fold: the whole decode+aggregate sequence
    algo_bytes = window_bytes + 16 * distinct  # ring bytes once + 16 B per distinct key
    achieved = algo_bytes / (span_ms * 1e-3) / 1e9
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": "c2_ip_hist: PERF_RECORD_SAMPLE ring decode + (cpu,pid,ip) histogram",
                   "cpus_per_gpu": args.cpus, "ring_mib": args.ring_mib, "sample_bytes": 40,
                   "records_per_gpu": n_rec, "distinct_keys_per_gpu": distinct,
                   "window_bytes_per_gpu": int(window_bytes),
                   "l2": "inputs (4 GiB/GPU) exceed L2; no flush needed",
                   "sharding": "by CPU id, 256 CPUs per rank; the decode needs no exchange (per-CPU outputs are "
                               "disjoint by key); at N > 1 every timed step ends with hbt_fetch_results_all, one "
                               "NCCL all-gather of every rank's per-CPU results",
                   "slow_tiles": slow, "reruns": reruns,
                   "max_over_ranks": {"kernel_ms": float(spread[0].item()),
                                      "host_queue_us_per_step": float(spread[1].item())}},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak,
                     "traffic": load_traffic("k_decode<IP_HISTOGRAM>") if args.agg == "hist" else None,
                     "kernel": "k_decode<IP_HISTOGRAM> + k_hist_aggregate" if args.agg == "hist" else "k_decode<NONE>",
                     "kernel_ms": span_ms, "decode_kernel_ms": k_ms, "fold_kernel_ms": f_ms,
                     "algorithmic_bytes": int(algo_bytes), "peak_source": peak_src,
                     "frac_of_nominal_8000": achieved / 8000.0},
        "gpu_launches": int(launches),
        "clocks": clocks,
    }
    if e2e:
        line["e2e"] = e2e
    if diagnostics:
        line["diagnostics"] = diagnostics
    if world == 1 and not args.no_cpu_baseline:
        os.sched_setaffinity(0, all_cpus)  # the CPU baseline gets every host core again
        cb = cpu_baseline_leg(args, gpu_sample)
        if cb:
            line["cpu_baseline"] = cb[0]
            if cb[1]:
                line["parity_check"] = cb[1]["result"]
                line["config"]["parity_check"] = cb[1]["checked"]
                assert cb[1]["result"] == "ok", "GPU histogram differs from the reference's on the sampled CPUs"
    print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
