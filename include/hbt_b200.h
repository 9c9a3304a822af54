/*
 * hbt_b200.h -- C-ABI of the B200-native perf_event decode/aggregate path.
 *
 * This is the drop-in boundary underneath dynolog/hbt's host C++ classes
 * (hbt::perf_event::CpuEventsGroup / PerCpuCountReader / GroupReadValues,
 * hbt::mon::Monitor, dynolog::PerfMonitor).  The reference has no FFI of its
 * own; every entry point below names the reference function (file:line under
 * the reference tree) whose work it replaces.  Host code in any language binds
 * these with plain pointers and sizes; no C++ or torch type crosses the ABI.
 *
 * Conventions
 *  - every call returns 0 / a non-negative count on success or a NEGATIVE errno
 *    (the convention of CpuEventsGroup::consume and its handlers,
 *    hbt/src/perf_event/CpuEventsGroup.h:1318-1326); no exception crosses.
 *  - one hbt_ctx per (calling thread, device); a ctx is not thread-safe, the
 *    same single-consumer rule the reference has for a CPU buffer.
 *  - the caller owns host buffers; the library owns device buffers unless an
 *    entry point is suffixed _dev, in which case the pointers are DEVICE
 *    pointers owned by the caller (e.g. torch tensors) and the call is
 *    asynchronous on the ctx stream.
 *  - there is no CPU fallback: without a usable CUDA device hbt_create fails
 *    with -ENODEV and nothing else can be called.
 *  - all integers little-endian, layouts identical to the Linux uapi
 *    (perf_event_header, read_format) and to GroupReadValues<TMode>::T
 *    (CpuEventsGroup.h:371-379): device layout == host layout == kernel layout.
 */
#ifndef HBT_B200_H_
#define HBT_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HBT_B200_ABI_VERSION 2

/* Limits taken from the reference: kMaxCpus (hbt/src/common/System.h:195),
 * CpuCountSampleGenerator::kMaxNumEvents (PerCpuCountSampleGenerator.h:44). */
#define HBT_MAX_CPUS 512
#define HBT_MAX_EVENTS 64

typedef struct hbt_ctx hbt_ctx;

/* ------------------------------------------------------------------------ */
/* Record model                                                             */
/* ------------------------------------------------------------------------ */

/* Index of the per-type record counters, in the order of the members
 * num_record_*_ of CpuEventsGroup (CpuEventsGroup.h:822-832). */
enum {
  HBT_CNT_LOST = 0,       /* += num_lost, not ++ (CpuEventsGroup.h:1381) */
  HBT_CNT_COMM = 1,
  HBT_CNT_EXIT = 2,
  HBT_CNT_THROTTLE = 3,
  HBT_CNT_UNTHROTTLE = 4,
  HBT_CNT_FORK = 5,
  HBT_CNT_SAMPLE = 6,
  HBT_CNT_READ = 7,
  HBT_CNT_AUX = 8,
  HBT_CNT_ITRACE_START = 9,
  HBT_CNT_SWITCH_CPU_WIDE = 10,
  HBT_NUM_COUNTERS = 11
};

/* Which record types have a (non-void) handler in the decode mode, i.e. which
 * `if constexpr (!std::is_void<TRecordX>)` bodies of consume() exist
 * (CpuEventsGroup.h:1372-1501).  A record of a type without handler is skipped
 * by header.size and only counted in the total. */
enum {
  HBT_H_LOST = 1u << 0,
  HBT_H_COMM = 1u << 1,
  HBT_H_EXIT = 1u << 2,
  HBT_H_FORK = 1u << 3,
  HBT_H_SAMPLE = 1u << 4,
  HBT_H_READ = 1u << 5,
  HBT_H_AUX = 1u << 6,
  HBT_H_ITRACE_START = 1u << 7,
  HBT_H_SWITCH_CPU_WIDE = 1u << 8
};

/* Handler sets of the reference's modes (CpuEventsGroup.h:66-291). */
#define HBT_MODE_SAMPLING (HBT_H_LOST | HBT_H_READ | HBT_H_SAMPLE)
#define HBT_MODE_CONTEXT_SWITCH \
  (HBT_H_LOST | HBT_H_COMM | HBT_H_EXIT | HBT_H_FORK | HBT_H_SWITCH_CPU_WIDE)
#define HBT_MODE_AUX_SPACE (HBT_H_AUX | HBT_H_ITRACE_START)
/* Mixed stream of BASELINE config 4: context-switch records plus SAMPLE. */
#define HBT_MODE_MIXED (HBT_MODE_CONTEXT_SWITCH | HBT_H_SAMPLE | HBT_H_READ)

/* Layout of PERF_RECORD_SAMPLE bodies. */
enum {
  /* mode::Sampling::Sample (CpuEventsGroup.h:268-281):
   * hdr tstamp nr time_enabled time_running value[nr]; size must equal
   * 40 + 8*nr (CpuEventsGroup.h:1435-1440) else -EPERM. */
  HBT_SAMPLE_GROUP = 0,
  /* sample_type = IP|TID|TIME|CPU in kernel order
   * (/usr/include/linux/perf_event.h:958-967):
   * hdr ip pid tid time cpu res [extra]; size must equal ip_sample_bytes.
   * No counterpart in the reference (it never sets PERF_SAMPLE_IP). */
  HBT_SAMPLE_IP = 1
};

/* What the SAMPLE handler does. */
enum {
  HBT_AGG_NONE = 0,        /* count only (handlers of CpuEventsGroupTest.cpp:58-88) */
  /* CpuCountSampleGenerator::handleRecordSample / handleRecordLost
   * (PerCpuCountSampleGenerator.h:110-202): rows {tstamp,duration,deltas[n]}. */
  HBT_AGG_COUNT_DELTAS = 1,
  /* (cpu,pid,ip) -> count table; needs HBT_SAMPLE_IP. */
  HBT_AGG_IP_HISTOGRAM = 2,
  /* CpuThreadSwitchGenerator's handlers (PerCpuThreadSwitchGenerator.h:228-412):
   * LOST / COMM / EXIT / FORK / SWITCH_CPU_WIDE records -> tagstack::Event stream
   * (hbt_switch_event).  handled_mask must be a subset of HBT_MODE_CONTEXT_SWITCH. */
  HBT_AGG_SWITCH_EVENTS = 3
};

/* tagstack::Event (hbt/src/tagstack/Event.h:28-227), same 16-byte layout.
 * Difference to the reference, by design: `tag` of ThreadCreation / ThreadDestruction
 * / SwitchIn events is the record's raw tid, not the virtual id PerThreadStats
 * hands out (that map is sequential host state shared by all CPUs; the host side
 * applies it: hbt_gpu::perf_event::ThreadSwitchVidMapper).  So that the host can do
 * this, COMM records -- which change that map but emit no reference Event -- appear
 * as pseudo-events of type HBT_SWEV_COMM (tag = tid), to be dropped after mapping. */
typedef struct hbt_switch_event {
  uint64_t tstamp;
  uint8_t type;          /* HBT_SWEV_* */
  uint8_t level;         /* desc->tid_level, or HBT_LEVEL_INVALID for WriteErrors* */
  uint16_t comp_unit_id; /* cpu_base + ctx-local cpu index */
  uint32_t tag;          /* tid, or HBT_TAG_NA */
} hbt_switch_event;

enum { /* Event::Type, Event.h:29-47 */
  HBT_SWEV_START = 0,
  HBT_SWEV_END = 1,
  HBT_SWEV_THREAD_CREATION = 2,
  HBT_SWEV_THREAD_DESTRUCTION = 3,
  HBT_SWEV_SWITCH_IN = 4,
  HBT_SWEV_SWITCH_OUT_PREEMPT = 5,
  HBT_SWEV_SWITCH_OUT_YIELD = 6,
  HBT_SWEV_WRITE_ERRORS_START = 7,
  HBT_SWEV_WRITE_ERRORS_END = 8,
  HBT_SWEV_COMM = 0x80 /* not a reference Event: see above */
};
#define HBT_TAG_NA 0xFFFFFFFFu    /* tagstack::kNA, TagStack.h:21 */
#define HBT_LEVEL_INVALID 0xFFu   /* tagstack::kInvalidLevel, TagStack.h:26 */

/* TSC -> kernel time parameters, TscConversionParams (System.h:175-179) and
 * CpuEventsGroup::kernelTimeFromTsc (CpuEventsGroup.h:1577-1591). */
typedef struct hbt_tsc_params {
  uint64_t time_zero;
  uint32_t time_mult;
  uint16_t time_shift;
  uint16_t pad_;
} hbt_tsc_params;

typedef struct hbt_decode_desc {
  uint32_t handled_mask;    /* HBT_H_* bits */
  uint32_t sample_layout;   /* HBT_SAMPLE_* */
  uint32_t aggregate;       /* HBT_AGG_* */
  uint32_t n_events;        /* HBT_AGG_COUNT_DELTAS: kNumEvents, 1..64 */
  uint32_t ip_sample_bytes; /* HBT_SAMPLE_IP: required header.size (>= 40, %8 == 0) */
  uint32_t max_records;     /* consume(max_num_records); 0xFFFFFFFF = drain */
  /* bit i set: event i has EventValueTransforms::TscToKernelTime
   * (PmuEvent.h:218-226, PerCpuCountSampleGenerator.h:171-173) */
  uint64_t tsc_transform_mask;
  hbt_tsc_params tsc;
  uint32_t tid_level;       /* HBT_AGG_SWITCH_EVENTS: tagstack level of thread tags (ctor arg tid_level) */
  uint32_t reserved_;       /* 0 */
} hbt_decode_desc;
#ifdef __cplusplus
static_assert(sizeof(hbt_decode_desc) == 56, "ABI 2 layout");
static_assert(sizeof(hbt_switch_event) == 16, "tagstack::Event layout");
#endif

/* Result of one decode call for one CPU: what consume() returned plus the
 * object state it mutated.  counters[] and num_unknown_count_interval are the
 * running totals since hbt_ring_reset_state, like the reference's members. */
typedef struct hbt_cpu_result {
  int64_t ret;       /* consume()'s return: #records, or -ENODATA (drained),
                        -EPERM (bad SAMPLE/READ size, tstamp regression), -EBADMSG
                        (header.size < 8: the reference would never return) */
  uint64_t n_records;/* records consumed by this call (also when ret < 0) */
  uint64_t new_tail; /* value consume() wrote to data_tail (CpuEventsGroup.h:1520) */
  uint64_t n_rows;   /* HBT_AGG_COUNT_DELTAS: rows emitted by this call;
                        HBT_AGG_SWITCH_EVENTS: events emitted by this call */
  uint64_t counters[HBT_NUM_COUNTERS];
  uint64_t num_unknown_count_interval; /* PerCpuCountSampleGenerator.h:152,230;
                        HBT_AGG_SWITCH_EVENTS: in_write_errors_ (0/1) after the call */
  uint64_t carry_tstamp; /* count_.tstamp after the call (UINT64_MAX = invalid);
                        HBT_AGG_SWITCH_EVENTS: last_tstamp_ (PerCpuThreadSwitchGenerator.h:437) */
} hbt_cpu_result;

/* One entry of the (cpu,pid,ip) table. */
typedef struct hbt_hist_entry {
  uint64_t ip;
  uint32_t pid;
  uint32_t cpu;
  uint64_t count;
} hbt_hist_entry;

/* ------------------------------------------------------------------------ */
/* Lifecycle                                                                */
/* ------------------------------------------------------------------------ */

/* ABI version of the loaded library. */
int hbt_abi_version(void);

/* Create a context on CUDA device `device`.  -ENODEV when there is no usable
 * sm_100 device, -EINVAL for max_cpus > HBT_MAX_CPUS or max_events > HBT_MAX_EVENTS.
 * Replaces nothing in the reference; it is the GPU analogue of constructing
 * the per-CPU generators (PerCpuBase.h:26-33). */
int hbt_create(hbt_ctx** out, int device, uint32_t max_cpus, uint32_t max_events);
void hbt_destroy(hbt_ctx* ctx);

/* Human-readable description of the last failure on this ctx (never NULL). */
const char* hbt_last_error(const hbt_ctx* ctx);

/* Run every later call of this ctx on `cuda_stream` (a cudaStream_t); NULL
 * restores the ctx's own stream.  A decode that is still pending is completed on
 * the stream it was queued on first (this call then blocks until it is done). */
int hbt_set_stream(hbt_ctx* ctx, void* cuda_stream);

/* Block until all work queued by this ctx is complete. */
int hbt_sync(hbt_ctx* ctx);

/* Number of kernels this ctx launched since creation (bench.py's gpu_launches). */
uint64_t hbt_launch_count(const hbt_ctx* ctx);

/* How a CPU's unread window is cut into segments (one CTA each): at most
 * segs_per_cpu segments of at least min_tiles_per_seg 16-KiB tiles.  0 = let the
 * library choose from the number of active CPUs and the SM count (the default).
 * Results never depend on it; it exists for benchmarks and tests that want to
 * exercise many segment seams.  No counterpart in the reference (one thread
 * walks a whole ring, CpuEventsGroup.h:1348). */
int hbt_set_geometry(hbt_ctx* ctx, uint32_t segs_per_cpu, uint32_t min_tiles_per_seg);

/* Which of the two decode kernels runs.  HBT_KERNEL_TILES (k_decode): a CTA per segment, tiles staged
 * in shared memory by TMA, a fast path for runs of equal-sized records -- the sampling streams.
 * HBT_KERNEL_WALK (k_walk): a warp per segment, every lane follows header.size through its own
 * kilobyte and the warp settles the seams -- streams of mixed record types (context-switch rings).
 * HBT_KERNEL_AUTO (the default) picks by the handled record types.  Results never depend on it (the
 * tests run both on the same input).  No counterpart in the reference. */
#define HBT_KERNEL_AUTO 0
#define HBT_KERNEL_TILES 1
#define HBT_KERNEL_WALK 2
int hbt_set_decode_kernel(hbt_ctx* ctx, int which);

/* CPU ids reported in hbt_hist_entry are cpu_base + the ctx-local index: a rank
 * that holds the shard [cpu_base, cpu_base + n_cpus) of a larger machine sets it
 * (sharding by CPU id, PerCpuBase.h:26 keeps one generator per CPU). */
int hbt_set_cpu_base(hbt_ctx* ctx, uint32_t cpu_base);

/* ------------------------------------------------------------------------ */
/* Ring ingest: CpuEventsGroup::onCpuDataBufferRead (CpuEventsGroup.h:1524-1570)
 * hands out the unread bytes of a CPU's ring as RbDataSlices (:44-63).       */
/* ------------------------------------------------------------------------ */

/* Library-owned device image of n_cpus rings of data_size bytes (power of two,
 * >= 4096: CpuEventsGroup.h:1065-1068,1077-1078).  Re-configuring frees the
 * previous images and resets all per-CPU state; with a different n_cpus it also
 * drops the histogram configuration (call hbt_hist_configure again). */
int hbt_ring_configure(hbt_ctx* ctx, uint32_t n_cpus, uint64_t data_size);

/* Copy the two slices of one OnRbReadCallback invocation into the device image
 * of `cpu` at ring position `tail & (data_size-1)` and set the unread window to
 * [tail, tail+n1+n2).  p2/n2 may be NULL/0.  n2 is clamped as the reference's
 * wrapped second slice must be (it reports data_size-n1, CpuEventsGroup.h:1561):
 * pass the true count head-tail-n1.  Host pointers; async H2D on the ctx stream
 * (pageable memory makes it synchronous). */
int hbt_submit_ring(hbt_ctx* ctx, uint32_t cpu, const void* p1, size_t n1,
                    const void* p2, size_t n2, uint64_t tail);

/* Use caller-owned DEVICE memory as the ring image of `cpu` (zero copy):
 * ring_base points at data_start (mmap_base + data_offset), data_size must be a
 * power of two, [tail, head) is the unread window (perf_event_mmap_page
 * data_tail / data_head, /usr/include/linux/perf_event.h:724-727).
 * ring_base must be 16-byte aligned. */
int hbt_submit_ring_dev(hbt_ctx* ctx, uint32_t cpu, const void* ring_base,
                        uint64_t data_size, uint64_t tail, uint64_t head);

/* The same for rings 0..n-1 in one call (all of one data_size). */
int hbt_submit_rings_dev(hbt_ctx* ctx, uint32_t n, const void* const* ring_bases,
                         uint64_t data_size, const uint64_t* tails, const uint64_t* heads);

/* Forget counters, carry (count_.tstamp, last_[]) and windows of all CPUs:
 * the state after enable() (CpuEventsGroup.h:1240-1280, enableImpl
 * PerCpuCountSampleGenerator.h:91-106).  seed_tstamp is the value enableImpl
 * takes from now(); UINT64_MAX seeds an invalid (re-seeding) interval.
 * The thread-switch generator's state goes back to its constructor values too:
 * last_tstamp_ = 0 (numeric_limits<TimeStamp>::min()), in_write_errors_ = false
 * (PerCpuThreadSwitchGenerator.h:437-438). */
int hbt_ring_reset_state(hbt_ctx* ctx, uint64_t seed_tstamp);

/* Restore the generator state of one CPU, e.g. after a daemon restart (the
 * reference has no checkpoint: this is the state its objects would have to carry,
 * SURVEY section 5): count_.tstamp and last_[n_events] of CpuCountSampleGenerator
 * (PerCpuCountSampleGenerator.h:222-230; UINT64_MAX = invalid, the next SAMPLE
 * re-seeds), the running per-type counters num_record_*_ (CpuEventsGroup.h:822-832,
 * HBT_NUM_COUNTERS words, NULL = keep) and num_unknown_count_interval_.  What
 * hbt_fetch_results (counters, carry_tstamp, num_unknown_count_interval) and
 * hbt_fetch_carry (last_[]) returned after the last decode can be fed back here.
 * Synchronous.  Completes a pending decode first. */
int hbt_ring_set_carry(hbt_ctx* ctx, uint32_t cpu, uint64_t carry_tstamp, const uint64_t* last,
                       uint32_t n_events, const uint64_t* counters, uint64_t num_unknown);
/* The same for CpuThreadSwitchGenerator's state: last_tstamp_ and in_write_errors_
 * (PerCpuThreadSwitchGenerator.h:437-438), as reported in hbt_cpu_result.carry_tstamp /
 * num_unknown_count_interval after an HBT_AGG_SWITCH_EVENTS decode. */
int hbt_ring_set_switch_state(hbt_ctx* ctx, uint32_t cpu, uint64_t last_tstamp,
                              uint32_t in_write_errors);

/* Back-pressure of the generators' output rings.  rows_per_cpu[c] = the rows CPU c's
 * sink can still take (its ring's free space / kWriteByteSize).  An
 * HBT_AGG_COUNT_DELTAS decode then behaves like the reference with a full ring and
 * no drop-if-full (PerCpuCountSampleGenerator.h:182-195; a handler returning < 0
 * leaves data_tail in FRONT of that record, CpuEventsGroup.h:1318-1322,1441-1444):
 * the SAMPLE whose row would not fit ends the CPU's consume() with ret = -ENOSPC,
 * new_tail at that SAMPLE, the record not counted, count_.tstamp invalid and
 * last_[] already holding its values -- submit the rest of the window again once
 * the sink has room and the SAMPLE re-seeds, exactly as it does in the reference.
 * The budgets stay in force until changed; NULL removes them.  n_cpus must equal
 * the configured CPU count. */
int hbt_set_rows_budget(hbt_ctx* ctx, const uint64_t* rows_per_cpu, uint32_t n_cpus);

/* ------------------------------------------------------------------------ */
/* Decode: CpuEventsGroup::consume (CpuEventsGroup.h:1326-1522) for every CPU
 * with a submitted window, plus the SAMPLE/LOST handlers selected by
 * desc->aggregate.                                                          */
/* ------------------------------------------------------------------------ */

/* Asynchronous on the ctx stream.  Returns 0 or -errno for invalid arguments /
 * CUDA failures; per-CPU outcomes are in hbt_fetch_results. */
int hbt_decode(hbt_ctx* ctx, const hbt_decode_desc* desc);

/* Synchronise and copy out one hbt_cpu_result per configured CPU (CPUs without
 * a window report ret = 0, n_records = 0).  Returns the number of CPUs. */
int hbt_fetch_results(hbt_ctx* ctx, hbt_cpu_result* out, uint32_t cap);

/* Diagnostics of the last fetched decode: how many times the batch had to be
 * re-run with corrected segment entries / clamped windows, and how many tiles
 * fell back to the one-thread sequential walk.  Both are 0 on well-formed streams. */
uint32_t hbt_decode_reruns(const hbt_ctx* ctx);
/* Segments decoded again by those re-runs.  A re-run is restricted to what the first
 * attempt got wrong: the one segment whose entry guess did not chain (every segment of
 * the CPU when its window has to be clamped, or under HBT_AGG_IP_HISTOGRAM); the other
 * CPUs' and segments' results stand. */
uint64_t hbt_decode_rerun_segments(const hbt_ctx* ctx);
uint64_t hbt_decode_slow_tiles(hbt_ctx* ctx);
/* Device time of the last decode kernel launch (CUDA events on the ctx stream
 * around that one launch; waits for it).  Negative when unavailable. */
float hbt_decode_kernel_ms(hbt_ctx* ctx);
/* Same for the last histogram fold kernel (k_hist_aggregate). */
float hbt_hist_fold_kernel_ms(hbt_ctx* ctx);

/* Rows {tstamp, duration, deltas[n_events]} of the last decode for `cpu`, in
 * stream order: the bytes CpuCountSampleGenerator writes to its output ring
 * (PerCpuCountSampleGenerator.h:69,182).  Returns the number of rows copied,
 * -ENOSPC if cap_rows is too small. */
int64_t hbt_fetch_rows(hbt_ctx* ctx, uint32_t cpu, uint64_t* rows, uint64_t cap_rows);

/* The rows of ALL CPUs of the last decode in one array: the per-segment slabs are packed on the
 * device (one launch) into [total rows][2 + n_events], CPU c at rows [offsets[c], offsets[c+1])
 * (offsets: host array of n_cpus + 1 entries, filled by the call), each CPU's rows in the order
 * its generator writes them to its output ring.  The pack kernel is QUEUED on the ctx's stream when
 * the call returns: consume the array in that stream's order, or call hbt_sync first when another
 * stream reads it.  hbt_fetch_rows_all then needs ONE device-to-host
 * copy for the whole decode (hbt_fetch_rows issues one per slab); the packed array stays valid
 * until the next decode.  Returns the total number of rows, -ENOSPC if cap_rows is too small. */
int hbt_rows_packed_dev(hbt_ctx* ctx, const uint64_t** rows, uint64_t* offsets);
int64_t hbt_fetch_rows_all(hbt_ctx* ctx, uint64_t* rows, uint64_t cap_rows, uint64_t* offsets);

/* Events of the last HBT_AGG_SWITCH_EVENTS decode for `cpu`, in the order
 * CpuThreadSwitchGenerator writes them to its output ring
 * (PerCpuThreadSwitchGenerator.h:228-412; kWriteByteSize = sizeof(Event), :186).
 * Returns the number copied, -ENOSPC if cap is too small. */
int64_t hbt_fetch_switch_events(hbt_ctx* ctx, uint32_t cpu, hbt_switch_event* out, uint64_t cap);
/* Device view: all CPUs' events back to back, CPU c at [offsets[c], offsets[c+1])
 * (offsets: host array of n_cpus + 1 entries, filled by the call). */
int hbt_switch_events_dev(hbt_ctx* ctx, const hbt_switch_event** events, uint64_t* offsets);
/* The per-CPU streams of that decode merged into one, in the order the reference's
 * binary tree of stream::Combinator nodes delivers them (hbt/src/tagstack/Stream.h:114-248,
 * makeBinaryTree :190-230): Event::happenedBefore = timestamp, then comp_unit_id
 * (Event.h:205-213).  Requires what the Combinator itself assumes: each CPU's stream is
 * timestamp-ordered.  The merge runs on the first call after a decode (log2(n_cpus)
 * launches, queued on the ctx's stream when the call returns: read the result in that stream's
 * order, or after hbt_sync) and the result stays valid until the next decode. */
int hbt_switch_events_merged_dev(hbt_ctx* ctx, const hbt_switch_event** merged, uint64_t* n_events);
int64_t hbt_fetch_switch_events_merged(hbt_ctx* ctx, hbt_switch_event* out, uint64_t cap);

/* Device view of the same rows for downstream kernels: rows of `cpu` live in
 * *n_slabs slabs; slab i has n_rows[i] rows at ptr[i].  Arrays must hold
 * hbt_rows_max_slabs() entries. */
int hbt_rows_slabs_dev(hbt_ctx* ctx, uint32_t cpu, const uint64_t** ptr,
                       uint64_t* n_rows, uint32_t* n_slabs);
uint32_t hbt_rows_max_slabs(const hbt_ctx* ctx);

/* last_[] of `cpu` after the last decode (PerCpuCountSampleGenerator.h:228). */
int hbt_fetch_carry(hbt_ctx* ctx, uint32_t cpu, uint64_t* last, uint32_t n_events);

/* (cpu,pid,ip) histogram.  Every CPU owns a set of append-only lists of 16-byte
 * entries {u64 ip; u32 pid; u32 count} (a key always goes to the list its hash
 * selects): a decode appends one entry per sample, then a fold kernel reduces each
 * list to distinct keys in shared memory, in place.  capacity_slots is the total
 * number of entries the lists can hold, split evenly between the configured CPUs;
 * per CPU it must cover the distinct keys accumulated so far PLUS the samples of
 * one decode, with headroom: lists hold 4096 entries and a key that repeats a few
 * hundred times within one decode fills its own list, so size for an average fill
 * of a quarter or less (capacity >= 4 x (distinct + samples per decode)).  A decode
 * that overflows a list makes hbt_fetch_results return -ENOSPC.  Counts are 32-bit and accumulate until hbt_hist_clear; hbt_decode
 * returns -EOVERFLOW rather than let more than 2^32-1 samples accumulate.
 * hbt_hist_clear only resets the list cursors. */
int hbt_hist_configure(hbt_ctx* ctx, uint64_t capacity_slots);
int hbt_hist_clear(hbt_ctx* ctx);
/* Number of distinct keys, then the entries (unordered). */
int64_t hbt_hist_size(hbt_ctx* ctx);
int64_t hbt_hist_fetch(hbt_ctx* ctx, hbt_hist_entry* out, uint64_t cap);
/* The entries of the ctx-local CPUs [first_cpu, first_cpu + n_cpus) only; out = NULL
 * returns their number. */
int64_t hbt_hist_fetch_cpus(hbt_ctx* ctx, uint32_t first_cpu, uint32_t n_cpus, hbt_hist_entry* out,
                            uint64_t cap);
/* The same entries in the 16-byte form the device keeps them in, CPU after CPU: the entries of
 * ctx-local CPU first_cpu + i are out[offsets[i] .. offsets[i + 1]) (offsets: host array of
 * n_cpus + 1 entries, filled by the call; unordered inside a CPU).  The cpu id travels in the
 * offset table instead of in every entry: a third fewer bytes over PCIe than hbt_hist_fetch_cpus,
 * which is what a collection loop that reads the table every interval pays for.  Returns the
 * number of entries, -ENOSPC if cap is too small; out = NULL only fills offsets. */
typedef struct hbt_hist_key_count {
  uint64_t ip;
  uint32_t pid;
  uint32_t count;
} hbt_hist_key_count;
int64_t hbt_hist_fetch_packed(hbt_ctx* ctx, uint32_t first_cpu, uint32_t n_cpus, hbt_hist_key_count* out,
                              uint64_t cap, uint64_t* offsets);
/* Device view for downstream kernels: list l = cpu * (*lists_per_cpu) + bucket
 * holds cursors[l] distinct entries at entries + l * (*list_capacity) * 16 bytes. */
int hbt_hist_lists_dev(hbt_ctx* ctx, const void** entries, const uint32_t** cursors,
                       uint32_t* lists_per_cpu, uint32_t* list_capacity);

/* ------------------------------------------------------------------------ */
/* Counter groups: PerCpuBase<T>::read (PerCpuBase.h:104-119) over a recorded
 * trace of GroupReadValues<TMode>::T, then GroupReadValues::getCount
 * (CpuEventsGroup.h:449-459) and PerfMonitor::log (PerfMonitor.cpp:52-67).  */
/* ------------------------------------------------------------------------ */

/* A derived ratio: double(getCount(num)) / double(getCount(den)), the
 * convention of PerfMonitor.cpp:56-57 applied to the event pairs that
 * BuiltinMetrics.cpp only describes (ipc :847-868, l3 misses/instr :508-532). */
typedef struct hbt_ratio {
  uint32_t num_event;
  uint32_t den_event;
} hbt_ratio;

/* Metric descriptors: which events make up a metric and in which order their values
 * appear in a group read (MetricDesc::event_refs_by_arch, hbt/src/perf_event/Metrics.h:45-179;
 * the registry is makeAvailableMetrics, BuiltinMetrics.cpp:470-1175).  The table inside the
 * library restates that registry (builtin = 1, pinned against it by
 * tests/golden/builtin_metrics.json) and adds two build-defined miss-rate groups over the
 * reference's generic hardware-cache events (builtin = 0).  ratios[] are NAMED derived
 * values over the descriptor's own event order; the reference defines none (no code divides
 * events), so they are build-defined in the convention of PerfMonitor.cpp:56-57. */
typedef struct hbt_metric_desc {
  const char* id;                 /* MetricDesc::id, e.g. "ipc" */
  const char* arch;               /* "any", or the CpuArch the event list is registered for */
  uint32_t n_events;
  const char* const* nicknames;   /* [n_events], index = position of the value (EventRef::nickname) */
  uint32_t n_ratios;
  const hbt_ratio* ratios;        /* [n_ratios] indices into nicknames */
  const char* const* ratio_names; /* [n_ratios] */
  uint32_t builtin;
  uint32_t index;                 /* position in the table */
  uint64_t tsc_transform_mask;    /* bit i: event i carries EventValueTransforms::TscToKernelTime */
} hbt_metric_desc;

uint32_t hbt_metric_count(void);
const hbt_metric_desc* hbt_metric_at(uint32_t index);
/* MetricDesc::getEventRefs(cpu_arch) (Metrics.h:162-180): the list registered for `arch` if
 * there is one, else the one registered for any architecture, else NULL.  arch may be NULL. */
const hbt_metric_desc* hbt_metric_find(const char* id, const char* arch);
/* MetricDesc::eventIdxFromNicknames (Metrics.h:136-151): -ENOENT if there is none. */
int hbt_metric_event_index(const hbt_metric_desc* desc, const char* nickname);

/* K4+K5 for one metric selected by descriptor: n_events and the ratio table are the
 * descriptor's (the index pairs live in constant memory on the device; nothing is uploaded).
 * ratio_out: [n_intervals][desc->n_ratios]. */
int hbt_counts_pipeline_metric_dev(hbt_ctx* ctx, const hbt_metric_desc* desc, const uint64_t* reads,
                                   uint64_t n_intervals, uint32_t n_cpus, uint64_t* sums,
                                   uint64_t* scaled, float* rates, double* ratio_out,
                                   uint64_t* deltas);

/* K4.  reads: [n_intervals][n_cpus][3+n_events] u64, one T per (interval,cpu).
 * sums:  [n_intervals][3+n_events] u64 = zero() then accum() over the CPUs
 * (CpuEventsGroup.h:508-522): word0 = n_events, time_enabled, time_running,
 * count[].  A T whose word0 != n_events is a failed read: the interval's word0
 * becomes 0 (PerCpuBase::read returns false, PerCpuCountReader::read nullopt). */
int hbt_counts_reduce_dev(hbt_ctx* ctx, const uint64_t* reads, uint64_t n_intervals,
                          uint32_t n_cpus, uint32_t n_events, uint64_t* sums);

/* K5.  From summed triples: scaled[i][e] = getCount(e); rates[i][e] =
 * float(double(scaled)*1e3/double(time_running)) (mips / mega_cycles_per_second);
 * ratios[i][r]; deltas[i][e] = scaled[i][e] - scaled[i-1][e] (u64 wrap; interval
 * 0 uses prev_scaled, or 0 when NULL): MetricSeries::diff convention
 * (dynolog/src/metric_frame/MetricSeries.h:221-227).  Any output may be NULL. */
int hbt_counts_scale_dev(hbt_ctx* ctx, const uint64_t* sums, uint64_t n_intervals,
                         uint32_t n_events, const hbt_ratio* ratios, uint32_t n_ratios,
                         const uint64_t* prev_scaled, uint64_t* scaled, float* rates,
                         double* ratio_out, uint64_t* deltas);

/* K4+K5 in one launch (single-GPU path: no reduction across ranks between). */
int hbt_counts_pipeline_dev(hbt_ctx* ctx, const uint64_t* reads, uint64_t n_intervals,
                            uint32_t n_cpus, uint32_t n_events, const hbt_ratio* ratios,
                            uint32_t n_ratios, uint64_t* sums, uint64_t* scaled,
                            float* rates, double* ratio_out, uint64_t* deltas);

/* Same with HOST buffers: H2D of reads, kernels, D2H of every non-NULL output,
 * then a sync.  This is what a PerCpuCountReader-shaped host adaptor calls. */
int hbt_counts_pipeline_host(hbt_ctx* ctx, const uint64_t* reads, uint64_t n_intervals,
                             uint32_t n_cpus, uint32_t n_events, const hbt_ratio* ratios,
                             uint32_t n_ratios, uint64_t* sums, uint64_t* scaled,
                             float* rates, double* ratio_out, uint64_t* deltas);

/* ------------------------------------------------------------------------ */
/* Metric tables: dynolog's MetricFrame (a set of MetricSeries rings sharing one
 * fixed-interval time axis, dynolog/src/metric_frame/MetricFrame.h:37-55,
 * MetricFrameTsUnit.cpp:14-132) kept in HBM, fed with the per-interval tables
 * the counter-group kernels produce, and the window statistics of
 * MetricSeriesSlice (MetricFrameBase.h:66-91 over MetricSeries.h:189-251).   */
/* ------------------------------------------------------------------------ */
typedef struct hbt_frame hbt_frame;
#define HBT_FRAME_MAX_LEN 8192u   /* samples a frame can hold (a window is staged in shared memory) */

enum {  /* MATCH_POLICY, MetricFrameTsUnitInterface.h:15-19 */
  HBT_MATCH_CLOSEST = 0,
  HBT_MATCH_PREV_CLOSEST = 1,
  HBT_MATCH_NEXT_CLOSEST = 2
};

/* MetricFrameRange (MetricFrameTsUnitInterface.h:21-29): offsets count from the
 * oldest sample held; times are steady_clock ticks (nanoseconds). */
typedef struct hbt_frame_range {
  uint64_t start_offset;
  uint64_t end_offset;      /* inclusive, like MetricFrameOffset */
  int64_t start_time_ns;
  int64_t end_time_ns;
} hbt_frame_range;

/* A MetricFrameVector of n_series rings of frame_len samples on ctx's device.
 * series_is_double[s] != 0 makes series s a MetricSeries<double>, else
 * MetricSeries<int64_t> (NULL: all int64).  interval_us is the fixed sampling
 * interval of the time axis (MetricFrameTsUnitFixInterval); 0 is refused like
 * the reference's constructor refuses it.  Destroy frames before their ctx. */
int hbt_frame_create(hbt_ctx* ctx, hbt_frame** out, uint32_t n_series, uint32_t frame_len,
                     const uint8_t* series_is_double, int64_t interval_us);
void hbt_frame_destroy(hbt_frame* frame);
uint64_t hbt_frame_length(const hbt_frame* frame);   /* MetricFrameBase::length() */

/* MetricFrameVector::addSamples for n_rows consecutive samples (oldest first):
 * row r holds one 8-byte value per series (int64 or the bits of a double);
 * last_time_ns is the time of the LAST row (with a fixed interval the axis keeps
 * nothing else).  _dev: rows is a DEVICE array with row_stride_words >= n_series
 * words per row -- e.g. the `scaled` / `deltas` / `ratio_out` table of
 * hbt_counts_pipeline_dev, appended without leaving the GPU; asynchronous on the
 * ctx stream.  The host variant copies the rows up and returns when done. */
int hbt_frame_add_samples_dev(hbt_frame* frame, const void* rows, uint64_t n_rows,
                              uint32_t row_stride_words, int64_t last_time_ns);
int hbt_frame_add_samples(hbt_frame* frame, const void* rows, uint64_t n_rows, int64_t last_time_ns);

/* MetricFrameBase::slice / MetricFrameTsUnitFixInterval::getRange
 * (MetricFrameTsUnit.cpp:49-99): the samples matching [start, end] under the two
 * policies.  -ENOENT where the reference returns std::nullopt. */
int hbt_frame_get_range(const hbt_frame* frame, int64_t start_ns, int64_t end_ns, int start_policy,
                        int end_policy, hbt_frame_range* out);

/* The statistics of MetricSeriesSlice for n_windows ranges x every series in one
 * pass; outputs are HOST arrays [n_windows][n_series], any may be NULL:
 *   diff      last - first of the window, raw 8 bytes (int64, or the bits of a double)
 *   rate_f64  rate<double>(period): double(diff) * us(period) / us(window duration)
 *   rate_i64  rate<int64_t>(period): diff * (period / duration) if period > duration
 *             else diff / (duration / period), on steady_clock ticks (MetricSeries.h:229-249)
 *   avg_f64   avg<double>(): left fold in sample order, divided by the sample count
 *   avg_i64   avg<int64_t>() (int64 series; 0 for double series)
 *   pct       percentile(percentage): the lround(percentage * (n - 1))-th order statistic, raw 8 bytes
 * A one-sample window has duration 0: the reference divides by zero there; here
 * rate_i64 is 0 and rate_f64 follows IEEE. */
int hbt_frame_query(hbt_frame* frame, const hbt_frame_range* ranges, uint32_t n_windows,
                    int64_t period_ns, float percentage, uint64_t* diff, double* rate_f64,
                    int64_t* rate_i64, double* avg_f64, int64_t* avg_i64, uint64_t* pct);

/* ------------------------------------------------------------------------ */
/* Slices: the interval join, per-stack-id frequencies and the interval x bin
 * duration matrix downstream of the decode (tagstack::IntervalSlicer,
 * hbt/src/tagstack/IntervalSlicer.cpp:52-216; mon::CuMonData::stackIdFreqs,
 * MonData::stackIdFreqs, hbt/src/mon/MonData.h:30-48,243-268,386-403;
 * binmap::TagStackIdBinner, :425-612).                                       */
/* ------------------------------------------------------------------------ */

/* tagstack::slice::Slice (hbt/src/tagstack/Slicer.h:30-84), same 32-byte layout. */
typedef struct hbt_slice {
  uint64_t tstamp;
  uint64_t duration;
  uint64_t stack_id;     /* TagStackId; UINT64_MAX = kInvalidTagStackId */
  uint8_t swin_type;     /* Slice::TransitionType: 0 NA, 1 Analysis, 2 ThreadPreempted, 3 ThreadYield, 4 PhaseChange */
  uint8_t swout_type;
  uint8_t pad_[6];
} hbt_slice;

/* mon::SliceFreq of one stack id (MonData.h:30-48). */
typedef struct hbt_slice_freq {
  uint64_t stack_id;
  uint64_t duration;       /* total duration */
  uint64_t num_obs;        /* times the slice appears */
  uint64_t num_intervals;  /* distinct intervals it appears in */
} hbt_slice_freq;

typedef struct hbt_slice_stats hbt_slice_stats;

/* Frequency table for up to max_stack_ids distinct stack ids on ctx's device
 * (-ENOSPC later if there are more).  Destroy it before its ctx. */
int hbt_slice_stats_create(hbt_ctx* ctx, hbt_slice_stats** out, uint64_t max_stack_ids);
void hbt_slice_stats_destroy(hbt_slice_stats* st);
int hbt_slice_stats_clear(hbt_slice_stats* st);

/* IntervalSlicer::consumeSlices for n_intervals intervals in one pass: for every
 * interval [tstamp, tstamp + duration) the slices makeIntervalIntersectingSlices
 * returns (trimmed copies of the boundary slices with TransitionType::Analysis when
 * `trim`; an empty list where the reference returns nullptr: nothing intersects,
 * or, with abort_on_missing_slices, the slices do not cover the interval without a
 * hole).  slices: DEVICE array of one compute unit's slices in time order (as a
 * Slicer writes them); offset = the consumer mark (slices already consumed).
 * intervals: DEVICE words, interval k = {tstamp, duration} at
 * intervals + k * interval_stride_words -- the packed delta rows of
 * hbt_rows_packed_dev qualify as they are (stride 2 + n_events), which is how the
 * reference's TraceCollector pairs count intervals with slices.  Intervals must be
 * in increasing time order.  out: DEVICE array of cap slices; out_offsets: DEVICE
 * array of n_intervals + 1 entries (interval k's slices are out[offsets[k] ..
 * offsets[k+1])).  *new_offset: the mark after the last interval that consumed
 * slices.  Returns the number of slices written, -ENOSPC if cap is too small,
 * -EINVAL for an interval of zero duration (HBT_ARG_CHECK_GT, :64) and -EDOM where
 * the reference throws because an interval's slices are longer than the interval
 * (:179-181; only possible without trimming). */
int64_t hbt_slices_join_dev(hbt_slice_stats* st, const hbt_slice* slices, uint64_t n_slices,
                            uint64_t offset, const uint64_t* intervals,
                            uint32_t interval_stride_words, uint64_t n_intervals, int trim,
                            int abort_on_missing_slices, hbt_slice* out, uint64_t cap,
                            uint64_t* out_offsets, uint64_t* new_offset);

/* CuMonData::stackIdFreqs of one compute unit's per-interval slices (the output of
 * the join), accumulated into the table like MonData::stackIdFreqs accumulates the
 * units (SliceFreq::accum): call once per compute unit. */
int hbt_slice_stats_add_unit_dev(hbt_slice_stats* st, const hbt_slice* pi_slices,
                                 const uint64_t* offsets, uint64_t n_intervals, uint64_t n_slices);
/* The accumulated frequencies in ascending stack-id order; out = NULL returns the count. */
int64_t hbt_slice_stats_fetch(hbt_slice_stats* st, hbt_slice_freq* out, uint64_t cap);

/* TagStackIdBinner (MonData.h:517-612).  prepare: every stack id of the table that
 * appears in >= min_distinct_intervals intervals with >= min_duration total gets a
 * bin of its own, the others share bin 0 (ignored_bid), and a zeroed n_rows x n_bins
 * matrix is made; returns n_bins and, if given, the stack ids (ascending) with their
 * bins.  Bins are numbered in ascending stack-id order (the reference numbers them
 * in the iteration order of an unordered_map, which is unspecified).  add_unit: one
 * compute unit's per-interval slices add their durations to matrix[row][bin], row =
 * rank of the interval among that unit's intervals with a non-zero total duration
 * (:563-606).  fetch: the matrix, row-major. */
int64_t hbt_slice_bins_prepare(hbt_slice_stats* st, uint64_t min_distinct_intervals,
                               uint64_t min_duration, uint64_t n_rows, uint64_t* stack_ids,
                               uint32_t* bids, uint64_t cap);
int hbt_slice_bins_add_unit_dev(hbt_slice_stats* st, const hbt_slice* pi_slices,
                                const uint64_t* offsets, uint64_t n_intervals);
int hbt_slice_bins_fetch(hbt_slice_stats* st, uint64_t* matrix, uint64_t cap_cells);

/* ------------------------------------------------------------------------ */
/* Multi-GPU merge (one process per GPU; the ranks of a box shard the CPUs).
 * The reference has no counterpart: it is a single host process.  What the
 * ranks exchange follows from its order of operations: PerCpuBase::read sums the
 * raw triples over the CPUs as u64 (PerCpuBase.h:104-119, CpuEventsGroup.h:514-522)
 * and getCount scales the SUM (:449-459), so partial sums are merged by an integer
 * all-reduce BEFORE the scaling; per-CPU outputs (results, rows, (cpu,..)-keyed
 * histogram) are disjoint by key and are only gathered.  The collectives are
 * NCCL's; libnccl.so.2 is bound at run time (the copy the process already holds,
 * e.g. torch's, else the system's), a single-GPU caller never loads it.       */
/* ------------------------------------------------------------------------ */
#define HBT_COMM_ID_BYTES 128   /* sizeof(ncclUniqueId) */

/* Rank 0 fills id_out (ncclGetUniqueId) and passes the bytes to the other ranks
 * over any host channel; then EVERY rank calls hbt_comm_init with them
 * (ncclCommInitRank, collective).  The communicator belongs to the ctx.
 * -ENOSYS when libnccl.so.2 cannot be loaded. */
int hbt_comm_unique_id(void* id_out);
int hbt_comm_init(hbt_ctx* ctx, const void* id, int n_ranks, int rank);
/* Or use a communicator the caller already has (an ncclComm_t made by the same
 * libnccl the process holds); it is not destroyed with the ctx. */
int hbt_comm_attach(hbt_ctx* ctx, void* nccl_comm, int n_ranks, int rank);
int hbt_comm_n_ranks(const hbt_ctx* ctx);   /* 0 = no communicator */

/* (i) sums [n_intervals][3+n_events] of this rank's CPU shard (hbt_counts_reduce_dev)
 * -> the sums over all ranks' CPUs, in place, on the ctx stream: ncclAllReduce(u64,
 * sum), then word0 = n_events if every shard read fine, else 0.  Feed the result to
 * hbt_counts_scale_dev. */
int hbt_counts_allreduce_dev(hbt_ctx* ctx, uint64_t* sums, uint64_t n_intervals, uint32_t n_events);

/* reduce -> all-reduce -> scale for a trace sharded by CPU: reads holds this rank's
 * n_cpus_local CPUs, every rank ends up with the outputs over ALL CPUs (same
 * outputs as hbt_counts_pipeline_dev over the unsharded trace; sums is required).
 * Works in interval chunks: the all-reduce of chunk k runs on the ctx's second
 * stream while chunk k+1 is reduced.  Collective: every rank calls it with the same
 * n_intervals / n_events. */
int hbt_counts_pipeline_sharded_dev(hbt_ctx* ctx, const uint64_t* reads, uint64_t n_intervals,
                                    uint32_t n_cpus_local, uint32_t n_events,
                                    const hbt_ratio* ratios, uint32_t n_ratios, uint64_t* sums,
                                    uint64_t* scaled, float* rates, double* ratio_out,
                                    uint64_t* deltas);

/* (ii) hbt_fetch_results on every rank at once: out_all[r * n_cpus + c] is CPU c of
 * rank r (ncclAllGather of the per-CPU results; all ranks configure the same
 * n_cpus).  Returns n_ranks * n_cpus. */
int hbt_fetch_results_all(hbt_ctx* ctx, hbt_cpu_result* out_all, uint32_t cap);

/* (ii) the (cpu,pid,ip) tables of all ranks on rank `root` (keys are disjoint by
 * cpu): root's `out` receives every rank's entries back to back, rank r's at
 * [sum(counts[0..r)), ...); counts_per_rank (n_ranks entries, may be NULL) is filled
 * on every rank.  Returns the number of entries in `out` on root, the rank's own
 * number elsewhere (out may be NULL there). */
int64_t hbt_hist_gather(hbt_ctx* ctx, int root, hbt_hist_entry* out, uint64_t cap,
                        uint64_t* counts_per_rank);

/* ------------------------------------------------------------------------ */
/* Allocation-free variant                                                  */
/* ------------------------------------------------------------------------ */

/* Largest n_intervals * n_events hbt_counts_deltas_static_dev accepts: its
 * working buffer is a fixed 4 MiB array inside the library image. */
#define HBT_STATIC_SCRATCH_WORDS (512u * 1024u)

/* No counterpart in the reference: this entry point exists for callers that must
 * not allocate device memory while they sample -- a monitor sharing a GPU whose
 * memory the watched job has all but exhausted (an allocation could fail, or push
 * the job over the edge), or a collection loop that cannot absorb the
 * device-wide stall a cudaMalloc/cudaFree causes.  It needs no hbt_ctx and calls
 * no cudaMalloc, cudaFree or any other allocating routine.
 *
 * Computes what the `deltas` output of hbt_counts_pipeline_dev holds:
 * deltas[i][e] = getCount(e) of interval i minus that of interval i-1 (interval 0
 * against 0), from reads[n_intervals][n_cpus][3+n_events].  reads and deltas are
 * DEVICE pointers on the CURRENT device; work is queued on `cuda_stream` (NULL =
 * the default stream) and is complete when that stream reaches it.
 *
 * Limits and rules:
 *  - n_intervals * n_events must not exceed HBT_STATIC_SCRATCH_WORDS; larger
 *    requests are refused with -E2BIG (split the trace by interval ranges and
 *    carry the last scaled values yourself, or use hbt_counts_pipeline_dev).
 *  - the first call on each device looks up the device address of the scratch
 *    array (cudaGetSymbolAddress); make that first call during start-up, before
 *    any stream capture begins on that device.
 *  - all calls on one device share the one scratch array: two calls whose GPU
 *    work overlaps (different streams, or different threads) corrupt each other.
 *    Serialise them on one stream.  The array keeps the previous call's values
 *    between calls; no call depends on them.
 * Returns 0, -EINVAL (bad arguments), -E2BIG (see above), -ENODEV (no current CUDA
 * device), -EIO (launch failure). */
int hbt_counts_deltas_static_dev(const uint64_t* reads, uint64_t n_intervals, uint32_t n_cpus,
                                 uint32_t n_events, uint64_t* deltas, void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* HBT_B200_H_ */
