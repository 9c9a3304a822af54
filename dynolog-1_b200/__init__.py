"""dynolog-1_b200: B200-native perf_event decode + aggregation for dynolog/hbt.

The product is the C-ABI shared library ``libhbt_b200.so`` (built in-tree from
``csrc/`` for sm_100a, declared in ``include/hbt_b200.h``).  This package is the
thin Python binding (ctypes) that tests and ``bench.py`` drive it through; the
C++ mirror of the reference's host classes lives in ``host/``.

There is no CPU fallback: importing works anywhere, but every compute entry
point needs the library and a B200; a missing library raises at first use.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REPO_ROOT = os.path.dirname(_HERE)
LIB_PATH = os.environ.get("HBT_B200_LIB") or os.path.join(_HERE, "libhbt_b200.so")  # env: tuning builds

# ---- constants of include/hbt_b200.h -------------------------------------------
ABI_VERSION = 2
MAX_CPUS = 512
MAX_EVENTS = 64
NUM_COUNTERS = 11
(CNT_LOST, CNT_COMM, CNT_EXIT, CNT_THROTTLE, CNT_UNTHROTTLE, CNT_FORK, CNT_SAMPLE, CNT_READ,
 CNT_AUX, CNT_ITRACE_START, CNT_SWITCH_CPU_WIDE) = range(11)
H_LOST, H_COMM, H_EXIT, H_FORK, H_SAMPLE, H_READ, H_AUX, H_ITRACE_START, H_SWITCH_CPU_WIDE = (
    1 << i for i in range(9))
MODE_SAMPLING = H_LOST | H_READ | H_SAMPLE
MODE_CONTEXT_SWITCH = H_LOST | H_COMM | H_EXIT | H_FORK | H_SWITCH_CPU_WIDE
MODE_AUX_SPACE = H_AUX | H_ITRACE_START
MODE_MIXED = MODE_CONTEXT_SWITCH | H_SAMPLE | H_READ
SAMPLE_GROUP, SAMPLE_IP = 0, 1
AGG_NONE, AGG_COUNT_DELTAS, AGG_IP_HISTOGRAM, AGG_SWITCH_EVENTS = 0, 1, 2, 3
KERNEL_AUTO, KERNEL_TILES, KERNEL_WALK = 0, 1, 2
# tagstack::Event::Type (Event.h:29-47) + the COMM pseudo-event of hbt_switch_event
(SWEV_START, SWEV_END, SWEV_THREAD_CREATION, SWEV_THREAD_DESTRUCTION, SWEV_SWITCH_IN,
 SWEV_SWITCH_OUT_PREEMPT, SWEV_SWITCH_OUT_YIELD, SWEV_WRITE_ERRORS_START,
 SWEV_WRITE_ERRORS_END) = range(9)
SWEV_COMM = 0x80
TAG_NA = 0xFFFFFFFF
LEVEL_INVALID = 0xFF
DRAIN = 0xFFFFFFFF
INVALID_TS = 0xFFFFFFFFFFFFFFFF


class TscParams(C.Structure):
    _fields_ = [("time_zero", C.c_uint64), ("time_mult", C.c_uint32),
                ("time_shift", C.c_uint16), ("pad_", C.c_uint16)]


class DecodeDesc(C.Structure):
    _fields_ = [("handled_mask", C.c_uint32), ("sample_layout", C.c_uint32),
                ("aggregate", C.c_uint32), ("n_events", C.c_uint32),
                ("ip_sample_bytes", C.c_uint32), ("max_records", C.c_uint32),
                ("tsc_transform_mask", C.c_uint64), ("tsc", TscParams),
                ("tid_level", C.c_uint32), ("reserved_", C.c_uint32)]


class CpuResult(C.Structure):
    _fields_ = [("ret", C.c_int64), ("n_records", C.c_uint64), ("new_tail", C.c_uint64),
                ("n_rows", C.c_uint64), ("counters", C.c_uint64 * NUM_COUNTERS),
                ("num_unknown_count_interval", C.c_uint64), ("carry_tstamp", C.c_uint64)]


class HistEntry(C.Structure):
    _fields_ = [("ip", C.c_uint64), ("pid", C.c_uint32), ("cpu", C.c_uint32),
                ("count", C.c_uint64)]


class Ratio(C.Structure):
    _fields_ = [("num_event", C.c_uint32), ("den_event", C.c_uint32)]


class FrameRange(C.Structure):
    _fields_ = [("start_offset", C.c_uint64), ("end_offset", C.c_uint64),
                ("start_time_ns", C.c_int64), ("end_time_ns", C.c_int64)]


FRAME_RANGE_DTYPE = np.dtype([("start_offset", "<u8"), ("end_offset", "<u8"),
                              ("start_time_ns", "<i8"), ("end_time_ns", "<i8")])
MATCH_CLOSEST, MATCH_PREV_CLOSEST, MATCH_NEXT_CLOSEST = 0, 1, 2
# tagstack::slice::Slice (Slicer.h:30-84) and mon::SliceFreq (MonData.h:30-48)
SLICE_DTYPE = np.dtype([("tstamp", "<u8"), ("duration", "<u8"), ("stack_id", "<u8"), ("swin_type", "u1"),
                        ("swout_type", "u1"), ("pad_", "u1", (6,))])
SLICE_FREQ_DTYPE = np.dtype([("stack_id", "<u8"), ("duration", "<u8"), ("num_obs", "<u8"), ("num_intervals", "<u8")])
assert SLICE_DTYPE.itemsize == 32
TT_NA, TT_ANALYSIS, TT_THREAD_PREEMPTED, TT_THREAD_YIELD, TT_PHASE_CHANGE = range(5)
FRAME_MAX_LEN = 8192


class MetricDesc(C.Structure):
    _fields_ = [("id", C.c_char_p), ("arch", C.c_char_p), ("n_events", C.c_uint32),
                ("nicknames", C.POINTER(C.c_char_p)), ("n_ratios", C.c_uint32),
                ("ratios", C.POINTER(Ratio)), ("ratio_names", C.POINTER(C.c_char_p)),
                ("builtin", C.c_uint32), ("index", C.c_uint32), ("tsc_transform_mask", C.c_uint64)]


HIST_DTYPE = np.dtype([("ip", "<u8"), ("pid", "<u4"), ("cpu", "<u4"), ("count", "<u8")])
RESULT_DTYPE = np.dtype([("ret", "<i8"), ("n_records", "<u8"), ("new_tail", "<u8"),
                         ("n_rows", "<u8"), ("counters", "<u8", (NUM_COUNTERS,)),
                         ("num_unknown_count_interval", "<u8"), ("carry_tstamp", "<u8")])
SWITCH_EVENT_DTYPE = np.dtype([("tstamp", "<u8"), ("type", "u1"), ("level", "u1"),
                               ("comp_unit_id", "<u2"), ("tag", "<u4")])
assert SWITCH_EVENT_DTYPE.itemsize == 16
HIST_PACKED_DTYPE = np.dtype([("ip", "<u8"), ("pid", "<u4"), ("count", "<u4")])   # hbt_hist_key_count
assert HIST_PACKED_DTYPE.itemsize == 16
assert HIST_DTYPE.itemsize == C.sizeof(HistEntry)
assert RESULT_DTYPE.itemsize == C.sizeof(CpuResult)

# Every symbol include/hbt_b200.h declares: (name, restype, argtypes)
_P = C.c_void_p
_U64P = C.POINTER(C.c_uint64)
EXPORTS = [
    ("hbt_abi_version", C.c_int, []),
    ("hbt_create", C.c_int, [C.POINTER(_P), C.c_int, C.c_uint32, C.c_uint32]),
    ("hbt_destroy", None, [_P]),
    ("hbt_last_error", C.c_char_p, [_P]),
    ("hbt_set_stream", C.c_int, [_P, _P]),
    ("hbt_sync", C.c_int, [_P]),
    ("hbt_launch_count", C.c_uint64, [_P]),
    ("hbt_set_geometry", C.c_int, [_P, C.c_uint32, C.c_uint32]),
    ("hbt_set_decode_kernel", C.c_int, [_P, C.c_int]),
    ("hbt_set_cpu_base", C.c_int, [_P, C.c_uint32]),
    ("hbt_ring_configure", C.c_int, [_P, C.c_uint32, C.c_uint64]),
    ("hbt_submit_ring", C.c_int, [_P, C.c_uint32, _P, C.c_size_t, _P, C.c_size_t, C.c_uint64]),
    ("hbt_submit_ring_dev", C.c_int, [_P, C.c_uint32, _P, C.c_uint64, C.c_uint64, C.c_uint64]),
    ("hbt_submit_rings_dev", C.c_int, [_P, C.c_uint32, _P, C.c_uint64, _P, _P]),
    ("hbt_ring_reset_state", C.c_int, [_P, C.c_uint64]),
    ("hbt_ring_set_carry", C.c_int, [_P, C.c_uint32, C.c_uint64, _P, C.c_uint32, _P, C.c_uint64]),
    ("hbt_ring_set_switch_state", C.c_int, [_P, C.c_uint32, C.c_uint64, C.c_uint32]),
    ("hbt_set_rows_budget", C.c_int, [_P, _P, C.c_uint32]),
    ("hbt_decode", C.c_int, [_P, C.POINTER(DecodeDesc)]),
    ("hbt_fetch_results", C.c_int, [_P, _P, C.c_uint32]),
    ("hbt_decode_reruns", C.c_uint32, [_P]),
    ("hbt_decode_rerun_segments", C.c_uint64, [_P]),
    ("hbt_decode_slow_tiles", C.c_uint64, [_P]),
    ("hbt_decode_kernel_ms", C.c_float, [_P]),
    ("hbt_hist_fold_kernel_ms", C.c_float, [_P]),
    ("hbt_fetch_rows", C.c_int64, [_P, C.c_uint32, _P, C.c_uint64]),
    ("hbt_rows_packed_dev", C.c_int, [_P, C.POINTER(_P), _U64P]),
    ("hbt_fetch_rows_all", C.c_int64, [_P, _P, C.c_uint64, _U64P]),
    ("hbt_fetch_switch_events", C.c_int64, [_P, C.c_uint32, _P, C.c_uint64]),
    ("hbt_switch_events_dev", C.c_int, [_P, C.POINTER(_P), _U64P]),
    ("hbt_switch_events_merged_dev", C.c_int, [_P, C.POINTER(_P), _U64P]),
    ("hbt_fetch_switch_events_merged", C.c_int64, [_P, _P, C.c_uint64]),
    ("hbt_rows_slabs_dev", C.c_int, [_P, C.c_uint32, C.POINTER(_P), _U64P, C.POINTER(C.c_uint32)]),
    ("hbt_rows_max_slabs", C.c_uint32, [_P]),
    ("hbt_fetch_carry", C.c_int, [_P, C.c_uint32, _P, C.c_uint32]),
    ("hbt_hist_configure", C.c_int, [_P, C.c_uint64]),
    ("hbt_hist_clear", C.c_int, [_P]),
    ("hbt_hist_size", C.c_int64, [_P]),
    ("hbt_hist_fetch", C.c_int64, [_P, _P, C.c_uint64]),
    ("hbt_hist_fetch_packed", C.c_int64, [_P, C.c_uint32, C.c_uint32, _P, C.c_uint64, _U64P]),
    ("hbt_hist_fetch_cpus", C.c_int64, [_P, C.c_uint32, C.c_uint32, _P, C.c_uint64]),
    ("hbt_hist_lists_dev", C.c_int,
     [_P, C.POINTER(_P), C.POINTER(_P), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    ("hbt_metric_count", C.c_uint32, []),
    ("hbt_metric_at", C.POINTER(MetricDesc), [C.c_uint32]),
    ("hbt_metric_find", C.POINTER(MetricDesc), [C.c_char_p, C.c_char_p]),
    ("hbt_metric_event_index", C.c_int, [C.POINTER(MetricDesc), C.c_char_p]),
    ("hbt_counts_pipeline_metric_dev", C.c_int,
     [_P, C.POINTER(MetricDesc), _P, C.c_uint64, C.c_uint32, _P, _P, _P, _P, _P]),
    ("hbt_counts_reduce_dev", C.c_int, [_P, _P, C.c_uint64, C.c_uint32, C.c_uint32, _P]),
    ("hbt_counts_scale_dev", C.c_int,
     [_P, _P, C.c_uint64, C.c_uint32, _P, C.c_uint32, _P, _P, _P, _P, _P]),
    ("hbt_counts_pipeline_dev", C.c_int,
     [_P, _P, C.c_uint64, C.c_uint32, C.c_uint32, _P, C.c_uint32, _P, _P, _P, _P, _P]),
    ("hbt_counts_pipeline_host", C.c_int,
     [_P, _P, C.c_uint64, C.c_uint32, C.c_uint32, _P, C.c_uint32, _P, _P, _P, _P, _P]),
    ("hbt_frame_create", C.c_int, [_P, C.POINTER(_P), C.c_uint32, C.c_uint32, _P, C.c_int64]),
    ("hbt_frame_destroy", None, [_P]),
    ("hbt_frame_length", C.c_uint64, [_P]),
    ("hbt_frame_add_samples_dev", C.c_int, [_P, _P, C.c_uint64, C.c_uint32, C.c_int64]),
    ("hbt_frame_add_samples", C.c_int, [_P, _P, C.c_uint64, C.c_int64]),
    ("hbt_frame_get_range", C.c_int, [_P, C.c_int64, C.c_int64, C.c_int, C.c_int, C.POINTER(FrameRange)]),
    ("hbt_frame_query", C.c_int,
     [_P, _P, C.c_uint32, C.c_int64, C.c_float, _P, _P, _P, _P, _P, _P]),
    ("hbt_slice_stats_create", C.c_int, [_P, C.POINTER(_P), C.c_uint64]),
    ("hbt_slice_stats_destroy", None, [_P]),
    ("hbt_slice_stats_clear", C.c_int, [_P]),
    ("hbt_slices_join_dev", C.c_int64,
     [_P, _P, C.c_uint64, C.c_uint64, _P, C.c_uint32, C.c_uint64, C.c_int, C.c_int, _P, C.c_uint64, _P, _U64P]),
    ("hbt_slice_stats_add_unit_dev", C.c_int, [_P, _P, _P, C.c_uint64, C.c_uint64]),
    ("hbt_slice_stats_fetch", C.c_int64, [_P, _P, C.c_uint64]),
    ("hbt_slice_bins_prepare", C.c_int64, [_P, C.c_uint64, C.c_uint64, C.c_uint64, _P, _P, C.c_uint64]),
    ("hbt_slice_bins_add_unit_dev", C.c_int, [_P, _P, _P, C.c_uint64]),
    ("hbt_slice_bins_fetch", C.c_int, [_P, _P, C.c_uint64]),
    ("hbt_comm_unique_id", C.c_int, [_P]),
    ("hbt_comm_init", C.c_int, [_P, _P, C.c_int, C.c_int]),
    ("hbt_comm_attach", C.c_int, [_P, _P, C.c_int, C.c_int]),
    ("hbt_comm_n_ranks", C.c_int, [_P]),
    ("hbt_counts_allreduce_dev", C.c_int, [_P, _P, C.c_uint64, C.c_uint32]),
    ("hbt_counts_pipeline_sharded_dev", C.c_int,
     [_P, _P, C.c_uint64, C.c_uint32, C.c_uint32, _P, C.c_uint32, _P, _P, _P, _P, _P]),
    ("hbt_fetch_results_all", C.c_int, [_P, _P, C.c_uint32]),
    ("hbt_hist_gather", C.c_int64, [_P, C.c_int, _P, C.c_uint64, _U64P]),
    ("hbt_counts_deltas_static_dev", C.c_int,
     [_P, C.c_uint64, C.c_uint32, C.c_uint32, _P, _P]),
]

_lib = None


def load_library(path: str = LIB_PATH) -> C.CDLL:
    """dlopen libhbt_b200.so and type every export.  Raises if it is missing:
    there is no other implementation to fall back to."""
    global _lib
    if _lib is not None and path == LIB_PATH:
        return _lib
    if not os.path.exists(path):
        raise RuntimeError(
            f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; "
            "g.build()'` (nvcc, sm_100a).  There is no CPU fallback.")
    lib = C.CDLL(path)
    for name, restype, argtypes in EXPORTS:
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = restype
        fn.argtypes = argtypes
    if lib.hbt_abi_version() != ABI_VERSION:
        raise RuntimeError("libhbt_b200.so ABI version mismatch")
    if path == LIB_PATH:
        _lib = lib
    return lib


class HbtError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"hbt error {code} ({os.strerror(-code) if code < 0 else code}): {msg}")
        self.code = code


def _ptr(x):
    """Raw pointer of a numpy array / torch tensor / int / None."""
    if x is None:
        return None
    if isinstance(x, int):
        return C.c_void_p(x)
    if isinstance(x, np.ndarray):
        assert x.flags["C_CONTIGUOUS"]
        return C.c_void_p(x.ctypes.data)
    if hasattr(x, "data_ptr"):  # torch tensor
        assert x.is_contiguous()
        return C.c_void_p(x.data_ptr())
    raise TypeError(type(x))


def make_ratios(pairs):
    arr = (Ratio * max(1, len(pairs)))()
    for i, (a, b) in enumerate(pairs):
        arr[i].num_event, arr[i].den_event = a, b
    return arr


class Context:
    """One hbt_ctx: a (thread, device) handle on the GPU decode/aggregate path."""

    def __init__(self, device: int = 0, max_cpus: int = MAX_CPUS, max_events: int = MAX_EVENTS):
        self.lib = load_library()
        h = C.c_void_p()
        rc = self.lib.hbt_create(C.byref(h), device, max_cpus, max_events)
        if rc != 0:
            raise HbtError(rc, "hbt_create failed (no usable sm_100 CUDA device?)")
        self.h = h
        self.kernel = KERNEL_AUTO
        self.device = device

    def close(self):
        if getattr(self, "h", None):
            self.lib.hbt_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc < 0:
            raise HbtError(int(rc), self.lib.hbt_last_error(self.h).decode())
        return rc

    # ---- plumbing
    def set_stream(self, stream_ptr):
        self._ck(self.lib.hbt_set_stream(self.h, C.c_void_p(stream_ptr) if stream_ptr else None))

    def use_torch_stream(self):
        """Queue this ctx's work on torch's current stream.  That must be a
        non-default stream (`torch.cuda.set_stream(torch.cuda.Stream())`): the
        default stream's handle is 0, which hbt_set_stream reads as "use the
        ctx's own stream", and events recorded on the default stream would then
        not see the ctx's kernels."""
        import torch
        h = torch.cuda.current_stream(self.device).cuda_stream
        if not h:
            raise RuntimeError("use_torch_stream() needs a non-default current torch stream")
        self.set_stream(h)

    def sync(self):
        self._ck(self.lib.hbt_sync(self.h))

    @property
    def launches(self) -> int:
        return int(self.lib.hbt_launch_count(self.h))

    def set_decode_kernel(self, which: int = 0):
        """KERNEL_AUTO / KERNEL_TILES (k_decode) / KERNEL_WALK (k_walk)."""
        self._ck(self.lib.hbt_set_decode_kernel(self.h, which))
        self.kernel = which

    def set_geometry(self, segs_per_cpu: int = 0, min_tiles_per_seg: int = 0):
        """Segments per CPU / tiles per segment of the decode (0 = automatic)."""
        self._ck(self.lib.hbt_set_geometry(self.h, segs_per_cpu, min_tiles_per_seg))

    def set_cpu_base(self, base: int):
        self._ck(self.lib.hbt_set_cpu_base(self.h, base))

    # ---- rings
    def ring_configure(self, n_cpus: int, data_size: int = 0):
        self._ck(self.lib.hbt_ring_configure(self.h, n_cpus, data_size))
        self.n_cpus = n_cpus

    def submit_ring(self, cpu: int, s1: np.ndarray, s2, tail: int):
        self._ck(self.lib.hbt_submit_ring(self.h, cpu, _ptr(s1), s1.nbytes,
                                          _ptr(s2) if s2 is not None and s2.nbytes else None,
                                          s2.nbytes if s2 is not None else 0, tail))

    def submit_ring_image(self, cpu: int, ring: np.ndarray, tail: int, head: int):
        """Host ring image + window -> the two RbDataSlices the reference's
        onCpuDataBufferRead would hand out (CpuEventsGroup.h:1524-1570)."""
        ds = ring.nbytes
        off = tail & (ds - 1)
        n = head - tail
        n1 = min(n, ds - off)
        flat = ring.view(np.uint8).reshape(-1)
        s1 = flat[off:off + n1]
        s2 = flat[0:n - n1] if n > n1 else None
        self.submit_ring(cpu, s1, s2, tail)

    def submit_ring_dev(self, cpu: int, ring_ptr, data_size: int, tail: int, head: int):
        self._ck(self.lib.hbt_submit_ring_dev(self.h, cpu, _ptr(ring_ptr), data_size, tail, head))

    def submit_rings_dev(self, bases: np.ndarray, data_size: int, tails: np.ndarray,
                         heads: np.ndarray):
        """bases/tails/heads: uint64 numpy arrays (device pointers, windows)."""
        self._ck(self.lib.hbt_submit_rings_dev(self.h, len(bases), _ptr(bases), data_size,
                                               _ptr(tails), _ptr(heads)))

    def reset_state(self, seed_tstamp: int = INVALID_TS):
        self._ck(self.lib.hbt_ring_reset_state(self.h, seed_tstamp))

    def set_carry(self, cpu: int, carry_tstamp: int, last=None, counters=None, num_unknown: int = 0):
        """Restore count_.tstamp / last_[] / num_record_*_ of one CPU (hbt_ring_set_carry)."""
        la = None if last is None else np.ascontiguousarray(last, dtype=np.uint64)
        ca = None if counters is None else np.ascontiguousarray(counters, dtype=np.uint64)
        assert ca is None or ca.shape[0] == NUM_COUNTERS
        self._ck(self.lib.hbt_ring_set_carry(self.h, cpu, carry_tstamp, _ptr(la),
                                             0 if la is None else la.shape[0], _ptr(ca), num_unknown))

    def set_rows_budget(self, rows_per_cpu=None):
        """Rows every CPU's sink can still take (hbt_set_rows_budget); None removes the budgets."""
        if rows_per_cpu is None:
            self._ck(self.lib.hbt_set_rows_budget(self.h, None, 0))
            return
        a = np.ascontiguousarray(rows_per_cpu, dtype=np.uint64)
        self._ck(self.lib.hbt_set_rows_budget(self.h, _ptr(a), a.shape[0]))

    def set_switch_state(self, cpu: int, last_tstamp: int, in_write_errors: int):
        self._ck(self.lib.hbt_ring_set_switch_state(self.h, cpu, last_tstamp, in_write_errors))

    # ---- decode
    def decode(self, desc: DecodeDesc):
        self._ck(self.lib.hbt_decode(self.h, C.byref(desc)))

    def fetch_results(self) -> np.ndarray:
        out = np.zeros(self.n_cpus, dtype=RESULT_DTYPE)
        self._ck(self.lib.hbt_fetch_results(self.h, _ptr(out), self.n_cpus))
        return out

    @property
    def reruns(self) -> int:
        return int(self.lib.hbt_decode_reruns(self.h))

    @property
    def rerun_segments(self) -> int:
        return int(self.lib.hbt_decode_rerun_segments(self.h))

    @property
    def decode_kernel_ms(self) -> float:
        return float(self.lib.hbt_decode_kernel_ms(self.h))

    @property
    def hist_fold_kernel_ms(self) -> float:
        return float(self.lib.hbt_hist_fold_kernel_ms(self.h))

    @property
    def slow_tiles(self) -> int:
        return int(self.lib.hbt_decode_slow_tiles(self.h))

    def fetch_rows(self, cpu: int, n_rows: int, n_events: int) -> np.ndarray:
        out = np.zeros((max(1, n_rows), 2 + n_events), dtype=np.uint64)
        n = self._ck(self.lib.hbt_fetch_rows(self.h, cpu, _ptr(out), out.shape[0]))
        return out[:n]

    def fetch_rows_all(self, n_events: int, out: np.ndarray | None = None, total_rows: int | None = None):
        """All CPUs' rows of the last decode with one device-to-host copy: (rows [total][2+n],
        offsets [n_cpus + 1]).  `out`: optional caller-owned (e.g. pinned) uint64 array."""
        off = np.zeros(self.n_cpus + 1, dtype=np.uint64)
        if out is None:
            if total_rows is None:
                base = C.c_void_p()
                self._ck(self.lib.hbt_rows_packed_dev(self.h, C.byref(base), off.ctypes.data_as(_U64P)))
                total_rows = int(off[self.n_cpus])
            out = np.zeros((max(1, total_rows), 2 + n_events), dtype=np.uint64)
        n = self._ck(self.lib.hbt_fetch_rows_all(self.h, _ptr(out), out.shape[0], off.ctypes.data_as(_U64P)))
        return out[:n], off

    def rows_packed_dev(self):
        """(device pointer, offsets[n_cpus + 1]) of all CPUs' rows of the last decode, packed back to back."""
        base = C.c_void_p()
        off = np.zeros(self.n_cpus + 1, dtype=np.uint64)
        self._ck(self.lib.hbt_rows_packed_dev(self.h, C.byref(base), off.ctypes.data_as(_U64P)))
        return base.value or 0, off

    def switch_events_merged_dev(self):
        """(device pointer, number of events) of all CPUs' events in Combinator order."""
        base, n = C.c_void_p(), C.c_uint64()
        self._ck(self.lib.hbt_switch_events_merged_dev(self.h, C.byref(base), C.byref(n)))
        return base.value or 0, int(n.value)

    def fetch_switch_events(self, cpu: int, n_events: int) -> np.ndarray:
        """Events of the last AGG_SWITCH_EVENTS decode for `cpu` (SWITCH_EVENT_DTYPE)."""
        out = np.zeros(max(1, n_events), dtype=SWITCH_EVENT_DTYPE)
        m = self._ck(self.lib.hbt_fetch_switch_events(self.h, cpu, _ptr(out), out.shape[0]))
        return out[:m]

    def switch_events_dev(self):
        """(device pointer, offsets[n_cpus + 1]) of all CPUs' events, back to back."""
        base = C.c_void_p()
        off = np.zeros(self.n_cpus + 1, dtype=np.uint64)
        self._ck(self.lib.hbt_switch_events_dev(self.h, C.byref(base), off.ctypes.data_as(_U64P)))
        return base.value or 0, off

    def fetch_switch_events_merged(self) -> np.ndarray:
        """All CPUs' events of the last AGG_SWITCH_EVENTS decode in Combinator order."""
        base, n = C.c_void_p(), C.c_uint64()
        self._ck(self.lib.hbt_switch_events_merged_dev(self.h, C.byref(base), C.byref(n)))
        out = np.zeros(max(1, n.value), dtype=SWITCH_EVENT_DTYPE)
        m = self._ck(self.lib.hbt_fetch_switch_events_merged(self.h, _ptr(out), out.shape[0]))
        return out[:m]

    def fetch_carry(self, cpu: int, n_events: int) -> np.ndarray:
        out = np.zeros(n_events, dtype=np.uint64)
        self._ck(self.lib.hbt_fetch_carry(self.h, cpu, _ptr(out), n_events))
        return out

    # ---- histogram
    def hist_configure(self, capacity_slots: int):
        self._ck(self.lib.hbt_hist_configure(self.h, capacity_slots))

    def hist_clear(self):
        self._ck(self.lib.hbt_hist_clear(self.h))

    def hist_size(self) -> int:
        return int(self._ck(self.lib.hbt_hist_size(self.h)))

    def hist_fetch(self, out: np.ndarray | None = None) -> np.ndarray:
        """Entries of the histogram.  `out`: optional caller-owned HIST_DTYPE array
        (e.g. a view of pinned memory, which makes the D2H copy several times
        faster); it must be large enough or -ENOSPC is raised."""
        if out is None:
            n = self.hist_size()
            out = np.zeros(max(1, n), dtype=HIST_DTYPE)
        m = self._ck(self.lib.hbt_hist_fetch(self.h, _ptr(out), out.shape[0]))
        return out[:m]

    def hist_fetch_packed(self, first_cpu: int = 0, n_cpus: int | None = None, out: np.ndarray | None = None):
        """(entries [HIST_PACKED_DTYPE, 16 bytes each], offsets [n_cpus + 1]): the entries of ctx-local CPU
        first_cpu + i are entries[offsets[i]:offsets[i + 1]].  `out`: optional caller-owned (pinned) array."""
        if n_cpus is None:
            n_cpus = self.n_cpus - first_cpu
        off = np.zeros(n_cpus + 1, dtype=np.uint64)
        if out is None:
            n = self._ck(self.lib.hbt_hist_fetch_packed(self.h, first_cpu, n_cpus, None, 0, off.ctypes.data_as(_U64P)))
            out = np.zeros(max(1, n), dtype=HIST_PACKED_DTYPE)
        m = self._ck(self.lib.hbt_hist_fetch_packed(self.h, first_cpu, n_cpus, _ptr(out), out.shape[0],
                                                    off.ctypes.data_as(_U64P)))
        return out[:m], off

    def hist_fetch_cpus(self, first_cpu: int, n_cpus: int) -> np.ndarray:
        """Entries of the ctx-local CPUs [first_cpu, first_cpu + n_cpus)."""
        n = self._ck(self.lib.hbt_hist_fetch_cpus(self.h, first_cpu, n_cpus, None, 0))
        out = np.zeros(max(1, n), dtype=HIST_DTYPE)
        m = self._ck(self.lib.hbt_hist_fetch_cpus(self.h, first_cpu, n_cpus, _ptr(out), out.shape[0]))
        return out[:m]

    def hist_lists_dev(self):
        """(entries pointer, cursors pointer, lists per CPU, list capacity) of the device lists."""
        ent, cur = C.c_void_p(), C.c_void_p()
        lpc, cap = C.c_uint32(), C.c_uint32()
        self._ck(self.lib.hbt_hist_lists_dev(self.h, C.byref(ent), C.byref(cur), C.byref(lpc), C.byref(cap)))
        return ent.value, cur.value, lpc.value, cap.value

    # ---- counter groups
    def counts_reduce_dev(self, reads, n_intervals, n_cpus, n_events, sums):
        self._ck(self.lib.hbt_counts_reduce_dev(self.h, _ptr(reads), n_intervals, n_cpus, n_events,
                                                _ptr(sums)))

    def counts_scale_dev(self, sums, n_intervals, n_events, ratios, prev_scaled=None, scaled=None,
                         rates=None, ratio_out=None, deltas=None):
        r = make_ratios(ratios)
        self._ck(self.lib.hbt_counts_scale_dev(self.h, _ptr(sums), n_intervals, n_events,
                                               C.cast(r, C.c_void_p), len(ratios),
                                               _ptr(prev_scaled), _ptr(scaled), _ptr(rates),
                                               _ptr(ratio_out), _ptr(deltas)))

    def counts_pipeline_dev(self, reads, n_intervals, n_cpus, n_events, ratios, sums=None,
                            scaled=None, rates=None, ratio_out=None, deltas=None):
        r = make_ratios(ratios)
        self._ck(self.lib.hbt_counts_pipeline_dev(self.h, _ptr(reads), n_intervals, n_cpus,
                                                  n_events, C.cast(r, C.c_void_p), len(ratios),
                                                  _ptr(sums), _ptr(scaled), _ptr(rates),
                                                  _ptr(ratio_out), _ptr(deltas)))

    def counts_pipeline_metric_dev(self, desc, reads, n_intervals, n_cpus, sums=None, scaled=None,
                                   rates=None, ratio_out=None, deltas=None):
        """K4+K5 for the metric descriptor `desc` (metric_find): events and named ratios are its."""
        self._ck(self.lib.hbt_counts_pipeline_metric_dev(self.h, desc, _ptr(reads), n_intervals, n_cpus,
                                                         _ptr(sums), _ptr(scaled), _ptr(rates),
                                                         _ptr(ratio_out), _ptr(deltas)))

    # ---- multi-GPU merge (NCCL)
    def comm_init(self, dist=None, group=None):
        """Give this ctx an NCCL communicator over the ranks of torch.distributed's (default)
        group: rank 0 draws the unique id, it travels as a broadcast, every rank joins."""
        if dist is None:
            import torch.distributed as dist
        import torch
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        ident = (C.c_ubyte * 128)()
        if rank == 0:
            rc = self.lib.hbt_comm_unique_id(ident)
            if rc < 0:
                raise HbtError(rc, "hbt_comm_unique_id (libnccl.so.2 not loadable?)")
        box = [bytes(ident)]
        dist.broadcast_object_list(box, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        buf = (C.c_ubyte * 128).from_buffer_copy(box[0])
        self._ck(self.lib.hbt_comm_init(self.h, buf, world, rank))
        self.n_ranks, self.rank = world, rank

    def comm_init_single(self):
        """A one-rank communicator (the collectives become local copies): single-GPU tests."""
        ident = (C.c_ubyte * 128)()
        rc = self.lib.hbt_comm_unique_id(ident)
        if rc < 0:
            raise HbtError(rc, "hbt_comm_unique_id (libnccl.so.2 not loadable?)")
        self._ck(self.lib.hbt_comm_init(self.h, ident, 1, 0))
        self.n_ranks, self.rank = 1, 0

    def counts_allreduce_dev(self, sums, n_intervals, n_events):
        self._ck(self.lib.hbt_counts_allreduce_dev(self.h, _ptr(sums), n_intervals, n_events))

    def counts_pipeline_sharded_dev(self, reads, n_intervals, n_cpus_local, n_events, ratios, sums,
                                    scaled=None, rates=None, ratio_out=None, deltas=None):
        r = make_ratios(ratios)
        self._ck(self.lib.hbt_counts_pipeline_sharded_dev(
            self.h, _ptr(reads), n_intervals, n_cpus_local, n_events, C.cast(r, C.c_void_p), len(ratios),
            _ptr(sums), _ptr(scaled), _ptr(rates), _ptr(ratio_out), _ptr(deltas)))

    def fetch_results_all(self) -> np.ndarray:
        """Per-CPU results of every rank: [n_ranks * n_cpus] (rank-major)."""
        out = np.zeros(self.n_cpus * self.n_ranks, dtype=RESULT_DTYPE)
        self._ck(self.lib.hbt_fetch_results_all(self.h, _ptr(out), out.shape[0]))
        return out

    def hist_gather(self, root: int = 0, out: np.ndarray | None = None, cap: int = 0):
        """(entries on root / None elsewhere, counts per rank)."""
        counts = np.zeros(self.n_ranks, dtype=np.uint64)
        if self.rank == root and out is None:
            out = np.zeros(max(1, cap), dtype=HIST_DTYPE)
        n = self._ck(self.lib.hbt_hist_gather(self.h, root, _ptr(out) if out is not None else None,
                                              0 if out is None else out.shape[0], counts.ctypes.data_as(_U64P)))
        return (out[:n] if self.rank == root else None), counts

    def counts_pipeline_host(self, reads: np.ndarray, ratios=()):
        """reads: [I][C][3+E] u64 on the host -> dict of host arrays."""
        I, Cc, W = reads.shape
        E = W - 3
        R = len(ratios)
        out = dict(sums=np.zeros((I, W), np.uint64), scaled=np.zeros((I, E), np.uint64),
                   rates=np.zeros((I, E), np.float32), ratios=np.zeros((I, max(R, 1)), np.float64),
                   deltas=np.zeros((I, E), np.uint64))
        r = make_ratios(ratios)
        self._ck(self.lib.hbt_counts_pipeline_host(
            self.h, _ptr(np.ascontiguousarray(reads)), I, Cc, E, C.cast(r, C.c_void_p), R,
            _ptr(out["sums"]), _ptr(out["scaled"]), _ptr(out["rates"]),
            _ptr(out["ratios"]) if R else None, _ptr(out["deltas"])))
        out["ratios"] = out["ratios"][:, :R]
        return out


class Frame:
    """hbt_frame: a MetricFrameVector kept on the device of `ctx` (dynolog/src/metric_frame)."""

    def __init__(self, ctx: "Context", is_double, frame_len: int, interval_us: int):
        self.ctx = ctx
        self.is_double = np.ascontiguousarray(is_double, dtype=np.uint8)
        self.n_series = len(self.is_double)
        h = C.c_void_p()
        ctx._ck(ctx.lib.hbt_frame_create(ctx.h, C.byref(h), self.n_series, frame_len, _ptr(self.is_double),
                                         interval_us))
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.ctx.lib.hbt_frame_destroy(self.h)
            self.h = None

    def __len__(self):
        return int(self.ctx.lib.hbt_frame_length(self.h))

    def add_samples(self, rows: np.ndarray, last_time_ns: int):
        rows = np.ascontiguousarray(rows)
        assert rows.ndim == 2 and rows.shape[1] == self.n_series and rows.itemsize == 8
        self.ctx._ck(self.ctx.lib.hbt_frame_add_samples(self.h, _ptr(rows), rows.shape[0], last_time_ns))

    def add_samples_dev(self, rows, n_rows: int, row_stride_words: int, last_time_ns: int):
        self.ctx._ck(self.ctx.lib.hbt_frame_add_samples_dev(self.h, _ptr(rows), n_rows, row_stride_words,
                                                            last_time_ns))

    def get_range(self, start_ns: int, end_ns: int, start_policy=MATCH_CLOSEST, end_policy=MATCH_CLOSEST):
        """MetricFrameBase::slice's range, or None where the reference returns std::nullopt."""
        r = FrameRange()
        rc = self.ctx.lib.hbt_frame_get_range(self.h, start_ns, end_ns, start_policy, end_policy, C.byref(r))
        if rc < 0:
            return None
        return (int(r.start_offset), int(r.end_offset), int(r.start_time_ns), int(r.end_time_ns))

    def query(self, ranges, period_ns: int, percentage: float):
        """ranges: list of get_range() tuples -> dict of [n_windows][n_series] arrays."""
        ra = np.zeros(len(ranges), dtype=FRAME_RANGE_DTYPE)
        for i, r in enumerate(ranges):
            ra[i] = r
        W, S = len(ranges), self.n_series
        out = dict(diff=np.zeros((W, S), np.uint64), rate_f64=np.zeros((W, S), np.float64),
                   rate_i64=np.zeros((W, S), np.int64), avg_f64=np.zeros((W, S), np.float64),
                   avg_i64=np.zeros((W, S), np.int64), pct=np.zeros((W, S), np.uint64))
        self.ctx._ck(self.ctx.lib.hbt_frame_query(self.h, _ptr(ra), W, period_ns, percentage, _ptr(out["diff"]),
                                                  _ptr(out["rate_f64"]), _ptr(out["rate_i64"]),
                                                  _ptr(out["avg_f64"]), _ptr(out["avg_i64"]), _ptr(out["pct"])))
        return out


class SliceStats:
    """hbt_slice_stats: the interval join, slice frequencies and bin matrix on the device of `ctx`."""

    def __init__(self, ctx: "Context", max_stack_ids: int):
        self.ctx = ctx
        h = C.c_void_p()
        ctx._ck(ctx.lib.hbt_slice_stats_create(ctx.h, C.byref(h), max_stack_ids))
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.ctx.lib.hbt_slice_stats_destroy(self.h)
            self.h = None

    def clear(self):
        self.ctx._ck(self.ctx.lib.hbt_slice_stats_clear(self.h))

    def join_dev(self, slices, n_slices, offset, intervals, stride_words, n_intervals, trim, abort_on_missing,
                 out, cap, out_offsets):
        """All device tensors/pointers; returns (slices written, new consumer mark)."""
        mark = C.c_uint64()
        n = self.ctx._ck(self.ctx.lib.hbt_slices_join_dev(self.h, _ptr(slices), n_slices, offset, _ptr(intervals),
                                                          stride_words, n_intervals, int(trim), int(abort_on_missing),
                                                          _ptr(out), cap, _ptr(out_offsets), C.byref(mark)))
        return int(n), int(mark.value)

    def add_unit_dev(self, pi_slices, offsets, n_intervals, n_slices):
        self.ctx._ck(self.ctx.lib.hbt_slice_stats_add_unit_dev(self.h, _ptr(pi_slices), _ptr(offsets), n_intervals,
                                                               n_slices))

    def fetch(self) -> np.ndarray:
        n = self.ctx._ck(self.ctx.lib.hbt_slice_stats_fetch(self.h, None, 0))
        out = np.zeros(max(1, n), dtype=SLICE_FREQ_DTYPE)
        m = self.ctx._ck(self.ctx.lib.hbt_slice_stats_fetch(self.h, _ptr(out), out.shape[0]))
        return out[:m]

    def bins_prepare(self, min_distinct_intervals, min_duration, n_rows):
        """-> (n_bins, stack ids ascending, their bins)."""
        n = self.ctx._ck(self.ctx.lib.hbt_slice_stats_fetch(self.h, None, 0))
        sids = np.zeros(max(1, n), dtype=np.uint64)
        bids = np.zeros(max(1, n), dtype=np.uint32)
        nb = self.ctx._ck(self.ctx.lib.hbt_slice_bins_prepare(self.h, min_distinct_intervals, min_duration, n_rows,
                                                              _ptr(sids), _ptr(bids), sids.shape[0]))
        self.n_rows, self.n_bins = n_rows, int(nb)
        return int(nb), sids[:n], bids[:n]

    def bins_add_unit_dev(self, pi_slices, offsets, n_intervals):
        self.ctx._ck(self.ctx.lib.hbt_slice_bins_add_unit_dev(self.h, _ptr(pi_slices), _ptr(offsets), n_intervals))

    def bins_fetch(self) -> np.ndarray:
        m = np.zeros((self.n_rows, self.n_bins), dtype=np.uint64)
        self.ctx._ck(self.ctx.lib.hbt_slice_bins_fetch(self.h, _ptr(m), m.size))
        return m


def metric_table():
    """The library's metric descriptors as dicts (id, arch, nicknames, ratios as
    (name, num, den) triples, builtin, tsc_transform_mask); no GPU needed."""
    lib = load_library()
    out = []
    for i in range(lib.hbt_metric_count()):
        d = lib.hbt_metric_at(i).contents
        out.append(dict(id=d.id.decode(), arch=d.arch.decode(),
                        nicknames=[d.nicknames[k].decode() for k in range(d.n_events)],
                        ratios=[(d.ratio_names[k].decode(), d.ratios[k].num_event, d.ratios[k].den_event)
                                for k in range(d.n_ratios)],
                        builtin=int(d.builtin), index=int(d.index), tsc_transform_mask=int(d.tsc_transform_mask)))
    return out


def metric_find(metric_id: str, arch: str | None = None):
    """hbt_metric_find: a pointer to the descriptor (pass it to counts_pipeline_metric_dev) or None."""
    p = load_library().hbt_metric_find(metric_id.encode(), arch.encode() if arch else None)
    return p if p else None


STATIC_SCRATCH_WORDS = 512 * 1024


def counts_deltas_static_dev(reads, n_intervals, n_cpus, n_events, deltas, stream_ptr=None) -> int:
    """hbt_counts_deltas_static_dev: the allocation-free entry point (no Context).
    reads / deltas are device tensors on the current device; returns the C return
    code (0 or a negative errno) instead of raising, so callers can test -E2BIG."""
    lib = load_library()
    return int(lib.hbt_counts_deltas_static_dev(_ptr(reads), n_intervals, n_cpus, n_events,
                                                _ptr(deltas),
                                                C.c_void_p(stream_ptr) if stream_ptr else None))


def sampling_desc(n_events: int, aggregate: int = AGG_COUNT_DELTAS, max_records: int = DRAIN,
                  tsc_mask: int = 0, tsc=(0, 1, 0), handled: int = MODE_SAMPLING) -> DecodeDesc:
    d = DecodeDesc()
    d.handled_mask = handled
    d.sample_layout = SAMPLE_GROUP
    d.aggregate = aggregate
    d.n_events = n_events
    d.ip_sample_bytes = 0
    d.max_records = max_records
    d.tsc_transform_mask = tsc_mask
    d.tsc = TscParams(tsc[0], tsc[1], tsc[2], 0)
    return d


def ip_desc(ip_sample_bytes: int = 40, aggregate: int = AGG_IP_HISTOGRAM,
            max_records: int = DRAIN) -> DecodeDesc:
    d = DecodeDesc()
    d.handled_mask = H_LOST | H_SAMPLE
    d.sample_layout = SAMPLE_IP
    d.aggregate = aggregate
    d.n_events = 0
    d.ip_sample_bytes = ip_sample_bytes
    d.max_records = max_records
    d.tsc = TscParams(0, 1, 0, 0)
    return d


def switch_desc(tid_level: int = 0, max_records: int = DRAIN,
                handled: int = MODE_CONTEXT_SWITCH) -> DecodeDesc:
    """mode::ContextSwitch + CpuThreadSwitchGenerator's handlers -> tagstack::Event stream."""
    d = DecodeDesc()
    d.handled_mask = handled
    d.sample_layout = SAMPLE_GROUP
    d.aggregate = AGG_SWITCH_EVENTS
    d.max_records = max_records
    d.tsc.time_mult = 1
    d.tid_level = tid_level
    return d


def count_desc(handled: int, max_records: int = DRAIN, sample_layout: int = SAMPLE_GROUP,
               ip_sample_bytes: int = 0) -> DecodeDesc:
    d = DecodeDesc()
    d.handled_mask = handled
    d.sample_layout = sample_layout
    d.aggregate = AGG_NONE
    d.ip_sample_bytes = ip_sample_bytes
    d.max_records = max_records
    d.tsc = TscParams(0, 1, 0, 0)
    return d
