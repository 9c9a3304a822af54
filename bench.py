#!/usr/bin/env python
"""bench.py -- records decoded+aggregated per second on synthetic perf_event traces.

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload ...]

Default workload = BASELINE.json configs[1]: PERF_RECORD_SAMPLE ring-buffer decode
+ (cpu,pid,ip) histogram, 256 CPUs x 16 MiB rings per GPU (IP|TID|TIME|CPU samples
of 40 B, pid ~ Zipf(1.1) over 4096, ip ~ Zipf(1.1) over 65536 addresses per pid).
One "step" = one pass of decode+aggregate over that batch.  Under torchrun every
rank holds its own 256-CPU shard (weak scaling, sharded by CPU id) and the
per-type record counters are summed with one NCCL all-reduce per step.

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
    """nvidia-smi clocks + throttle reasons during the timed region.  The process is started
    ahead of the region (its start-up takes the driver's locks for a moment and would otherwise
    stall the first timed step of every rank on the box); stop(t0, t1) keeps the samples that
    were read inside [t0, t1]."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.samples = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "20"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append((time.time(), line.strip()))

    def stop(self, t0=None, t1=None):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        lines = [l for ts, l in self.samples if t0 is None or t0 <= ts <= t1 + 0.02]
        if not lines:  # a region shorter than the sampling period: the samples closest to it
            lines = [l for _, l in sorted(self.samples, key=lambda x: abs(x[0] - t1))[:3]]
        for s in lines:
            parts = [x.strip() for x in s.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for n, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


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
    if switch:
        imgs, tails, heads, n_rec = make_switch_rings(args.cpus, ring_bytes, rank * args.cpus, dev)
    elif mixed:
        imgs, tails, heads, n_rec = make_mixed_rings(args.cpus, ring_bytes, rank * args.cpus, dev, E)
    else:
        imgs, tails, heads, n_rec = make_group_rings(args.cpus, ring_bytes, rank * args.cpus, dev, E)
    window_bytes = sum(h - t for t, h in zip(tails, heads))
    ctx = hbt.Context(local, max_cpus=args.cpus)
    ctx.use_torch_stream()
    ctx.ring_configure(args.cpus, 0)
    desc = hbt.sampling_desc(E, handled=hbt.MODE_MIXED if mixed else hbt.MODE_SAMPLING)
    if switch:
        desc = hbt.switch_desc(tid_level=1)
        ctx.set_cpu_base(rank * args.cpus)
    if args.agg == "none":  # diagnostic: boundary discovery + per-type counters only
        desc = hbt.count_desc(hbt.MODE_CONTEXT_SWITCH if switch else hbt.MODE_MIXED if mixed
                              else hbt.MODE_SAMPLING)
    bases_a = np.array([imgs[c].data_ptr() for c in range(args.cpus)], dtype=np.uint64)
    tails_a, heads_a = np.array(tails, dtype=np.uint64), np.array(heads, dtype=np.uint64)
    kms = []

    def step():
        # every step decodes the same trace again: start from the state after enable(),
        # or the second pass would (correctly) report a timestamp regression
        ctx.reset_state(hbt.INVALID_TS)
        ctx.submit_rings_dev(bases_a, ring_bytes, tails_a, heads_a)
        ctx.decode(desc)
        step.res = ctx.fetch_results()
        kms.append(ctx.decode_kernel_ms)

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()  # ahead of the warm-up: see ClockSampler
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
    clocks = sampler.stop(wall0, time.time()) if rank == 0 else None
    if rank != 0:
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
        step()
        torch.cuda.synchronize()
        m0.record()
        ctx._ck(ctx.lib.hbt_switch_events_merged_dev(ctx.h, C.byref(base), C.byref(n_ev)))
        m1.record()
        torch.cuda.synchronize()
        merge_ms = m0.elapsed_time(m1)
        return {"metric": METRIC, "value": n_rec * world / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world,
                "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_step,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64",
                "data": "synthetic",
                "config": {"workload": "c4_switch: mode::ContextSwitch records (32-48 B) -> tagstack::Event stream",
                           "cpus_per_gpu": args.cpus, "ring_mib": args.ring_mib, "records_per_gpu": n_rec,
                           "events_per_gpu": n_rows, "window_bytes_per_gpu": int(window_bytes),
                           "merge_ms": merge_ms, "merge_levels": max(0, (args.cpus - 1).bit_length()),
                           "l2": "inputs exceed L2; no flush needed", "slow_tiles": ctx.slow_tiles,
                           "reruns": ctx.reruns},
                "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                             "traffic": None, "kernel": "k_decode<SWITCH_EVENTS> + k_sw_count/scan/emit",
                             "kernel_ms": ms_step, "decode_kernel_ms": k_ms, "algorithmic_bytes": int(algo),
                             "peak_source": peak_src},
                "gpu_launches": int(launches), "clocks": clocks,
                **({"cpu_baseline": sw_cb} if sw_cb else {})}
    ach = algo / (k_ms * 1e-3) / 1e9
    cb = None
    if world == 1 and not args.no_cpu_baseline and not mixed and not switch and args.agg != "none":
        cb = cpu_baseline_deltas(args, E)
    line = {"metric": METRIC, "value": n_rec * world / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64",
            "data": "synthetic",
            "config": {"workload": (f"c4_mixed: SAMPLE/SWITCH/COMM/MMAP/LOST/FORK/EXIT stream, per-type counters "
                                    f"+ per-interval deltas ({E} events)") if mixed else
                                   f"c3_deltas: group SAMPLE ({40 + 8 * E} B, {E} events) decode + per-interval deltas",
                       "cpus_per_gpu": args.cpus, "ring_mib": args.ring_mib, "records_per_gpu": n_rec,
                       "rows_per_gpu": n_rows, "window_bytes_per_gpu": int(window_bytes),
                       "l2": "inputs exceed L2; no flush needed", "slow_tiles": ctx.slow_tiles,
                       "reruns": ctx.reruns},
            "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                         "traffic": (load_traffic(f"k_decode<COUNT_DELTAS>@{E}ev") if not mixed else None),
                         "kernel": "k_decode<COUNT_DELTAS>",
                         "kernel_ms": k_ms, "algorithmic_bytes": int(algo), "peak_source": peak_src},
            "gpu_launches": int(launches), "clocks": clocks}
    if cb:
        line["cpu_baseline"] = cb
    return line


def bench_counts(args, rank, world, local, dev, dist):
    """Diagnostic workload (config 5 shape): counter-group reduce over CPUs, NCCL sum of the
    integer triples across ranks, then scaling + rates + IPC ratio + interval deltas."""
    from dynolog_1_b200 import sharding
    E, C, I = args.events, args.cpus, args.intervals
    reads = tg.counts_trace(tg.stream_seed(5, rank), I, C, E, device=dev, multiplexed=True)
    ctx = hbt.Context(local, max_cpus=max(C, 1))
    ctx.use_torch_stream()
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
            ctx.counts_reduce_dev(reads, I, C, E, sums)
            sharding.allreduce_partial_sums(sums, E)
            ctx.counts_scale_dev(sums, I, E, ratios, None, scaled, rates, ratio, deltas)

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()  # ahead of the warm-up: see ClockSampler
    for _ in range(max(args.warmup, 3)):
        step()
    ctx.sync()
    l0 = ctx.launches
    wall0 = time.time()
    ms_step = timed_steps(step, args.steps, dist, dev)
    launches = ctx.launches - l0
    clocks = sampler.stop(wall0, time.time()) if rank == 0 else None
    if rank != 0:
        return None
    peak, peak_src = load_peaks()
    algo = I * C * W * 8 + I * (W * 8 + E * 8 + E * 4 + 8 + E * 8)
    ach = algo / (ms_step * 1e-3) / 1e9
    return {"metric": "counter-group reads decoded+aggregated/sec", "value": I * C * world / (ms_step * 1e-3),
            "unit": "reads/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u64/f64", "data": "synthetic",
            "config": {"workload": f"c5_counts: {C} CPUs x {E} events x {I} intervals per GPU, "
                                   "reduce -> [NCCL sum] -> scale/rates/ratio/deltas",
                       "l2": f"input {I * C * W * 8 / 2**20:.0f} MiB per step exceeds L2" if I * C * W * 8 > 2**27
                             else "input smaller than L2: not a valid HBM measurement"},
            "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                         "traffic": load_traffic("k_counts_reduce"), "kernel": "k_counts_reduce (+ scale, diff)",
                         "kernel_ms": ms_step, "algorithmic_bytes": int(algo), "peak_source": peak_src,
                         "note": "timed as the whole step (2-3 launches), not one kernel"},
            "gpu_launches": int(launches), "clocks": clocks}


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
    sample_cpus = min(args.cpus, 4)  # ~4 us per record (a /proc lookup per new thread): keep it short
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


def cpu_baseline_leg(args):
    """Bounded CPU baseline beside the GPU number (rank 0, N=1 only)."""
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
        t = lib.hbtref_ip_hist(h, rings.ctypes.data, ring_bytes, tails_a.ctypes.data,
                               heads_a.ctypes.data, sample_cpus, 40, hbt.DRAIN, cores, res)
        best = t if best is None else min(best, t)
    lib.hbtref_hist_free(h)
    # one thread, as the reference runs it (one collecting thread), on 4 of the rings
    one = min(4, sample_cpus)
    h1 = lib.hbtref_hist_new()
    t1 = lib.hbtref_ip_hist(h1, rings.ctypes.data, ring_bytes, tails_a.ctypes.data, heads_a.ctypes.data,
                            one, 40, hbt.DRAIN, 1, res)
    lib.hbtref_hist_free(h1)
    return {"value": n_rec / best, "unit": UNIT, "cores": cores, "kind": "reference",
            "sample": f"{sample_cpus} of {args.cpus} rings x {args.ring_mib} MiB, best of 2; "
                      "reference consume() + std::unordered_map handler",
            "single_thread_value": (n_rec // sample_cpus) * one / t1 if t1 > 0 else None}


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
    ap.add_argument("--workload", default="c2_ip_hist", choices=["c2_ip_hist", "c3_deltas", "c4_mixed", "c4_switch", "c5_counts"],
                    help="c2_ip_hist is the headline (BASELINE configs[1]); the others are diagnostics")
    ap.add_argument("--events", type=int, default=8)
    ap.add_argument("--intervals", type=int, default=100000)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-groups", type=int, default=8,
                    help="CPU groups (contexts) the end-to-end leg pipelines over PCIe")
    ap.add_argument("--no-cpu-baseline", action="store_true")
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

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()  # long before the timed region: see ClockSampler
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
    desc = hbt.ip_desc(40, aggregate=hbt.AGG_IP_HISTOGRAM if args.agg == "hist" else hbt.AGG_NONE)
    counters_dev = torch.zeros(hbt.NUM_COUNTERS + 1, dtype=torch.int64, device=dev)

    bases_a = np.array([imgs[c].data_ptr() for c in range(args.cpus)], dtype=np.uint64)
    tails_a = np.array(tails, dtype=np.uint64)
    heads_a = np.array(heads, dtype=np.uint64)

    def step():
        if args.agg == "hist":
            ctx.hist_clear()
        ctx.submit_rings_dev(bases_a, ring_bytes, tails_a, heads_a)
        ctx.decode(desc)

    def step_exchange():
        # metric-table merge across the GPUs of the box: per-type record counters
        if dist is not None:
            from dynolog_1_b200 import sharding
            sharding.allreduce_record_counters(counters_dev)

    # ---- warm-up + validation -------------------------------------------------
    for it in range(max(args.warmup, 3)):
        step()
        res = ctx.fetch_results()
        # counters are running totals since reset_state, like the reference's members
        assert int(res["counters"][:, hbt.CNT_SAMPLE].sum()) == n_rec * (it + 1)
    total = int(res["n_records"].sum())
    assert total == n_rec, (total, n_rec)
    distinct = ctx.hist_size() if args.agg == "hist" else 0
    slow, reruns = ctx.slow_tiles, ctx.reruns

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
        # the decode path shards by CPU with no exchange step: no collective in here
        h0 = time.perf_counter()
        step()
        host_us.append((time.perf_counter() - h0) * 1e6)  # host time to queue one step
        # completing the decode (re-run check) is part of the step; it syncs the stream
        last_res = ctx.fetch_results()
        kernel_ms.append(ctx.decode_kernel_ms)
        if args.agg == "hist":
            fold_ms.append(ctx.hist_fold_kernel_ms)
    e1.record()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    ms_total = e0.elapsed_time(e1)
    launches = ctx.launches - launches0
    clocks = sampler.stop(wall0, time.time()) if rank == 0 else None
    # the metric-table merge across the GPUs of the box, once, outside the timed steps: the
    # per-type record counters of every rank's CPUs summed over NCCL (and checked)
    counters_dev.zero_()
    counters_dev[:hbt.NUM_COUNTERS] = torch.from_numpy(
        last_res["counters"].sum(axis=0).astype(np.int64)).to(dev)
    step_exchange()
    merged_samples = int(counters_dev[hbt.CNT_SAMPLE].item())
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
        groups = []
        ent_dt = hbt.HIST_DTYPE.itemsize
        for g in range(G):
            lo, hi = g * per, min(args.cpus, (g + 1) * per)
            if lo >= hi:
                break
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
            groups.append((cg, lo, hi, pinned.numpy().view(hbt.HIST_DTYPE), (pinned, sg)))

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
                n_ent += len(cg.hist_fetch(out=buf))
                trace.append(time.perf_counter())
            if os.environ.get("HBT_BENCH_TRACE"):
                print("e2e trace (ms since issue):",
                      " ".join(f"{(x - trace[0]) * 1e3:.1f}" for x in trace[1:]), file=sys.stderr)
            return np.concatenate(rs), n_ent

        e2e_step()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(max(1, min(args.steps, 3))):
            r, n_ent = e2e_step()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / max(1, min(args.steps, 3))
        tt = torch.tensor([dt], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt.item())
        assert int(r["n_records"].sum()) == n_rec
        assert n_ent == distinct, (n_ent, distinct)
        e2e = {"value": n_rec * world / dt, "unit": UNIT, "h2d_bytes_per_step": int(window_bytes),
               "d2h_bytes_per_step": int(n_ent * ent_dt + r.nbytes), "ms_per_step": dt * 1e3,
               "pipeline": f"{len(groups)} CPU groups, one hbt_ctx + stream each: H2D of group g+1 "
                           "overlaps decode + fold + D2H of group g",
               "host_cpus_bound": len(numa) if numa else None}
        for cg, *_ in groups:
            cg.close()

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
                   "sharding": "by CPU id, 256 CPUs per rank, no collective in the decode path; per-type counters "
                               "merged over NCCL once after the timed steps",
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
    if world == 1 and not args.no_cpu_baseline:
        os.sched_setaffinity(0, all_cpus)  # the CPU baseline gets every host core again
        cb = cpu_baseline_leg(args)
        if cb:
            line["cpu_baseline"] = cb
    print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
