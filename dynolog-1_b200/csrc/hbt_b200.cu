// libhbt_b200.so: the C-ABI of include/hbt_b200.h over the sm_100a kernels in
// counts.cuh / decode.cuh.  Host orchestration only: buffers, segment tables,
// launches, and the (rare) exact re-run loop.  There is no CPU implementation of
// any entry point: without a CUDA device hbt_create fails and nothing else runs.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <cerrno>
#include <cstdarg>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/hbt_b200.h"
#include "counts.cuh"
#include "ctx_internal.h"
#include "decode.cuh"
#include "walk.cuh"
#include "metric_table.h"

using namespace hbt;

namespace {

constexpr int kMaxSegsPerCpu = 256;
constexpr int kMaxRatios = 64;
constexpr int kMaxCommChunks = 8;  // interval chunks of hbt_counts_pipeline_sharded_dev

template <class T>
struct DevBuf {
  T* p = nullptr;
  size_t n = 0;  // capacity in elements
  cudaError_t reserve(size_t want) {
    if (want <= n) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr;
    n = 0;
    cudaError_t e = cudaMalloc(&p, want * sizeof(T));
    if (e == cudaSuccess) n = want;
    return e;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    n = 0;
  }
};

template <class T>
struct PinBuf {
  T* p = nullptr;
  size_t n = 0;
  cudaError_t reserve(size_t want) {
    if (want <= n) return cudaSuccess;
    if (p) cudaFreeHost(p);
    p = nullptr;
    n = 0;
    cudaError_t e = cudaMallocHost(&p, want * sizeof(T));
    if (e == cudaSuccess) n = want;
    return e;
  }
  void release() {
    if (p) cudaFreeHost(p);
    p = nullptr;
    n = 0;
  }
};

void comm_release(hbt_ctx* ctx);  // destroys an owned NCCL communicator (end of this file)

struct RowSlab {  // one run of rows of the last COUNT_DELTAS decode
  const uint64_t* src;
  uint64_t dst_row;  // first row in the packed array
  uint64_t n_rows;
};

struct Window {
  const uint8_t* base = nullptr;
  uint64_t data_size = 0;
  uint64_t tail = 0, head = 0;
  bool active = false;
};

}  // namespace

struct hbt_ctx {
  int device = 0;
  int sm_count = 148;
  uint32_t max_cpus = 0, max_events = 0;
  cudaStream_t own_stream = nullptr;
  cudaStream_t stream = nullptr;
  std::string err;
  uint64_t launches = 0;

  // rings
  uint32_t n_cpus = 0;
  uint64_t data_size = 0;
  DevBuf<uint8_t> ring_store;  // library-owned images
  std::vector<Window> win;
  uint32_t cpu_base = 0;

  // generator state (double-buffered: [cur] is committed)
  DevBuf<CpuState> state[2];
  int cur = 0;

  // decode scratch
  DevBuf<RingDesc> d_rings;
  DevBuf<SegDesc> d_segs;
  DevBuf<SegResult> d_results;
  DevBuf<hbt_cpu_result> d_out;
  DevBuf<uint8_t> d_seam;
  DevBuf<uint64_t> d_forced;
  DevBuf<uint64_t> d_clamp_head, d_locate_from, d_locate_skip;
  DevBuf<uint32_t> d_locate_mode;
  DevBuf<uint64_t> d_rows_budget;  // hbt_set_rows_budget
  bool has_rows_budget = false;
  DevBuf<int64_t> d_clamp_ret;
  DevBuf<uint32_t> d_flags;
  DevBuf<uint64_t> d_rows;
  // hbt_rows_compact_dev: the slabs of the last decode packed into one array
  DevBuf<uint64_t> d_rows_packed;
  PinBuf<RowSlab> h_slabs;
  std::vector<uint64_t> rows_offsets;  // [n_cpus + 1], rows
  bool rows_packed_valid = false;
  PinBuf<RingDesc> h_rings;
  PinBuf<SegDesc> h_segs;
  PinBuf<SegResult> h_results;
  PinBuf<hbt_cpu_result> h_out;
  PinBuf<uint8_t> h_seam;
  PinBuf<uint64_t> h_clamp_head;
  PinBuf<int64_t> h_clamp_ret;
  PinBuf<uint32_t> h_flags;
  uint32_t n_segs = 0;
  // a re-run decodes only the segments of h_run_list (n_run = 0: every segment) and finalises
  // only the CPUs that are not settled yet (d_cpu_done)
  DevBuf<uint32_t> d_run_list;
  PinBuf<uint32_t> h_run_list;
  uint32_t n_run = 0;
  DevBuf<uint8_t> d_cpu_done;
  DevBuf<uint32_t> d_seg_flag;  // per segment: its exit is published (k_walk's look-back)
  PinBuf<uint8_t> h_cpu_done;
  uint64_t rerun_segments = 0;  // segments decoded again by the re-runs of the last decode
  uint32_t geo_segs_per_cpu = 0, geo_min_tiles = 0;  // hbt_set_geometry; 0 = automatic
  int kernel_choice = HBT_KERNEL_AUTO;                // hbt_set_decode_kernel
  bool walk = false;                                  // this decode runs k_walk (set by build_tables)
  bool walk_attr_set[4] = {false, false, false, false};
  bool tiled = true;
  bool attr_set[4] = {false, false, false, false};
  bool pending = false;  // a decode was launched and not yet fetched
  bool have_results = false;
  hbt_decode_desc desc{};
  uint32_t reruns = 0;
  cudaEvent_t ev_k0 = nullptr, ev_k1 = nullptr;  // around the decode kernel
  cudaEvent_t ev_f0 = nullptr, ev_f1 = nullptr;  // around the histogram fold kernel
  bool fold_timed = false;

  // HBT_AGG_SWITCH_EVENTS: generator state (last_tstamp_, in_write_errors_), pass-2 scratch, events
  DevBuf<uint64_t> sw_state_ts;
  DevBuf<uint32_t> sw_state_err;
  DevBuf<SwSeg> d_sw;
  DevBuf<uint64_t> d_sw_off;
  PinBuf<uint64_t> h_sw_off;
  DevBuf<hbt_switch_event> sw_events;
  DevBuf<hbt_switch_event> sw_merge[2];       // ping-pong buffers of the Combinator tree levels
  DevBuf<uint32_t> d_merge_prefix;
  DevBuf<uint64_t> d_merge_cuts;  // per merge tile: where its first diagonal cuts the pair's first run
  PinBuf<uint32_t> h_merge_prefix;
  const hbt_switch_event* sw_merged = nullptr;  // result of the last merge, nullptr = not merged yet
  bool sw_valid = false;  // sw_events / h_sw_off describe the last fetched decode

  // histogram: per-(cpu,bucket) lists (see HistLists)
  DevBuf<ulonglong2> hist_entries;
  DevBuf<uint32_t> hist_cursors, hist_cursors_snap;
  DevBuf<uint32_t> hist_overflow;
  uint32_t hist_buckets = 0, hist_cap = 0;
  uint32_t hist_n_cpus = 0;        // the n_cpus the lists were sized for
  bool hist_pending_fold = false;  // the last decode appended raw samples not yet folded
  bool agg_attr_set = false;
  bool counts_attr_set = false;
  uint64_t hist_samples = 0;       // upper bound of samples added since the last clear
  DevBuf<hbt_hist_entry> hist_out;
  DevBuf<unsigned long long> hist_n;
  // hbt_hist_fetch_packed: per-list offsets inside its CPU, per-CPU totals + their prefix, the packed entries
  DevBuf<uint32_t> hist_list_off;
  DevBuf<unsigned long long> hist_cpu_total;
  PinBuf<unsigned long long> h_hist_off;
  DevBuf<ulonglong2> hist_packed;

  // multi-GPU merge (NCCL communicator over the ranks of the box)
  void* comm = nullptr;       // ncclComm_t
  bool comm_owned = false;
  int n_ranks = 1, rank = 0;
  cudaStream_t comm_stream = nullptr;
  cudaEvent_t ev_chunk_reduced[kMaxCommChunks] = {}, ev_chunk_merged[kMaxCommChunks] = {};
  DevBuf<hbt_cpu_result> d_out_all;
  PinBuf<hbt_cpu_result> h_out_all;
  DevBuf<unsigned long long> d_hist_counts;
  PinBuf<unsigned long long> h_hist_counts;

  // counts scratch
  const hbt_ratio* metric_ratios_dev = nullptr;  // c_metric_ratios on this device
  DevBuf<hbt_ratio> d_ratios;
  DevBuf<uint64_t> c_reads, c_sums, c_scaled, c_deltas;
  DevBuf<float> c_rates;
  DevBuf<double> c_ratio;
};

namespace {

int fail(hbt_ctx* c, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  if (c) c->err = buf;
  return code;
}

int cuda_fail(hbt_ctx* c, cudaError_t e, const char* what) {
  return fail(c, e == cudaErrorMemoryAllocation ? -ENOMEM : -EIO, "%s: %s", what,
              cudaGetErrorString(e));
}

#define CU(call)                                       \
  do {                                                 \
    cudaError_t e__ = (call);                          \
    if (e__ != cudaSuccess) return cuda_fail(ctx, e__, #call); \
  } while (0)

bool is_pow2(uint64_t x) { return x && !(x & (x - 1)); }

size_t decode_smem_bytes(int agg) {
  size_t s = (size_t)kStages * kTile + ((sizeof(Smem) + 15) & ~size_t(15));
  if (agg == HBT_AGG_IP_HISTOGRAM) s += sizeof(HistCache);
  if (agg == HBT_AGG_COUNT_DELTAS) s += sizeof(RowDescs);
  return s;
}

int validate_desc(hbt_ctx* ctx, const hbt_decode_desc* d) {
  if (!d) return fail(ctx, -EINVAL, "desc is NULL");
  if (d->sample_layout > HBT_SAMPLE_IP) return fail(ctx, -EINVAL, "bad sample_layout");
  if (d->aggregate > HBT_AGG_SWITCH_EVENTS) return fail(ctx, -EINVAL, "bad aggregate");
  if (d->aggregate == HBT_AGG_SWITCH_EVENTS) {
    if (d->handled_mask == 0 || (d->handled_mask & ~(uint32_t)HBT_MODE_CONTEXT_SWITCH))
      return fail(ctx, -EINVAL, "SWITCH_EVENTS handles mode::ContextSwitch records only");
    if (d->tid_level > 0xFF) return fail(ctx, -EINVAL, "tid_level %u", d->tid_level);
  }
  if (d->aggregate == HBT_AGG_COUNT_DELTAS) {
    if (d->sample_layout != HBT_SAMPLE_GROUP || !(d->handled_mask & HBT_H_SAMPLE))
      return fail(ctx, -EINVAL, "COUNT_DELTAS needs handled group SAMPLE records");
    if (d->n_events == 0 || d->n_events > ctx->max_events)
      return fail(ctx, -EINVAL, "n_events %u outside 1..%u", d->n_events, ctx->max_events);
    if (d->tsc_transform_mask && d->tsc.time_shift >= 64)
      return fail(ctx, -EINVAL, "time_shift >= 64");
  }
  if (d->aggregate == HBT_AGG_IP_HISTOGRAM) {
    if (d->sample_layout != HBT_SAMPLE_IP || !(d->handled_mask & HBT_H_SAMPLE))
      return fail(ctx, -EINVAL, "IP_HISTOGRAM needs handled IP SAMPLE records");
    if (ctx->hist_buckets == 0 || ctx->hist_n_cpus != ctx->n_cpus)
      return fail(ctx, -EINVAL, "hbt_hist_configure first (after the rings are configured)");
  }
  if (d->sample_layout == HBT_SAMPLE_IP &&
      (d->ip_sample_bytes < 40 || (d->ip_sample_bytes & 7) || d->ip_sample_bytes > 65528))
    return fail(ctx, -EINVAL, "bad ip_sample_bytes %u", d->ip_sample_bytes);
  return 0;
}

// Which kernel decodes: k_walk for streams of mixed record types (anything but a sampling-only
// handler set), where k_decode's uniform fast path has nothing to hold on to; k_decode otherwise.
static bool use_walk(const hbt_ctx* ctx) {
  const hbt_decode_desc& d = ctx->desc;
  if (d.aggregate == HBT_AGG_IP_HISTOGRAM) return false;
  if (ctx->kernel_choice != HBT_KERNEL_AUTO) return ctx->kernel_choice == HBT_KERNEL_WALK;
  // measured (profiles/r2_notes.md): k_walk wins on streams of mixed record types, k_decode's uniform fast
  // path on runs of equal-sized samples
  return (d.handled_mask & (HBT_H_COMM | HBT_H_EXIT | HBT_H_FORK | HBT_H_SWITCH_CPU_WIDE | HBT_H_AUX |
                            HBT_H_ITRACE_START)) != 0;
}

// Build ring / segment tables for the current windows (+ clamps) into pinned
// memory and upload them.
int build_tables(hbt_ctx* ctx, const uint64_t* clamp_head, const int64_t* clamp_ret,
                 const std::vector<SegDesc>* keep_forced) {
  const hbt_decode_desc& d = ctx->desc;
  const uint32_t C = ctx->n_cpus;
  CU(ctx->h_rings.reserve(C));
  CU(ctx->d_rings.reserve(C));

  uint32_t active = 0;
  bool any_small = false;
  for (uint32_t c = 0; c < C; ++c) {
    const Window& w = ctx->win[c];
    if (w.active && w.head > w.tail) {
      ++active;
      if (w.data_size < 2ull * kTile) any_small = true;
    }
  }
  ctx->tiled = !any_small;
  ctx->walk = ctx->tiled && use_walk(ctx);

  // segments per CPU: enough CTAs to fill the machine a few times, at least
  // 8 tiles (128 KiB) per segment.  k_walk gives a segment to a WARP: many more, shorter segments.
  const uint64_t target_ctas = (uint64_t)ctx->sm_count * (ctx->walk ? 96 : 8);
  uint32_t segs_per_cpu = active ? (uint32_t)((target_ctas + active - 1) / active) : 1;
  // the histogram's per-segment cache of hot keys wants short segments
  if (d.aggregate == HBT_AGG_IP_HISTOGRAM && segs_per_cpu < 16) segs_per_cpu = 16;
  if (ctx->geo_segs_per_cpu) segs_per_cpu = ctx->geo_segs_per_cpu;  // hbt_set_geometry
  if (segs_per_cpu > kMaxSegsPerCpu) segs_per_cpu = kMaxSegsPerCpu;
  if (segs_per_cpu < 1) segs_per_cpu = 1;

  // record size when the stream is expected to be one run of equal-sized SAMPLE records (0: none)
  uint32_t hint_size = 0;
  if (d.sample_layout == HBT_SAMPLE_IP && (d.handled_mask & HBT_H_SAMPLE)) hint_size = d.ip_sample_bytes;
  else if (d.aggregate == HBT_AGG_COUNT_DELTAS) hint_size = 40 + 8 * d.n_events;
  if (hint_size > (uint32_t)kTile / 2) hint_size = 0;
  uint64_t min_tiles_per_seg = ctx->walk ? 2 : 8;
  if (ctx->geo_min_tiles) min_tiles_per_seg = ctx->geo_min_tiles;
  std::vector<SegDesc> segs;
  segs.reserve((size_t)active * segs_per_cpu + 1);  // no regrowth while the tables are built
  const uint64_t row_words = 2 + (uint64_t)d.n_events;
  uint64_t rows_total_words = 0;
  const bool sw = d.aggregate == HBT_AGG_SWITCH_EVENTS;  // 16-byte proto-events in the row slabs
  const bool with_rows = sw || d.aggregate == HBT_AGG_COUNT_DELTAS;
  // first pass: segment geometry and row slab sizes
  for (uint32_t c = 0; c < C; ++c) {
    const Window& w = ctx->win[c];
    RingDesc& r = ctx->h_rings.p[c];
    memset(&r, 0, sizeof(r));
    r.base = w.base;
    r.mask = w.data_size ? w.data_size - 1 : 0;
    r.tail = w.tail;
    r.true_head = w.head;
    r.head = w.head;
    r.override_ret = 0;
    if (clamp_head && clamp_head[c] != kNone) {
      r.head = clamp_head[c] < w.head ? clamp_head[c] : w.head;
      r.override_ret = clamp_ret[c];
    }
    r.seg_begin = (uint32_t)segs.size();
    r.n_seg = 0;
    if (!w.active || w.head <= w.tail) continue;  // geometry follows the TRUE window
    if (!ctx->tiled) {
      SegDesc s{};
      s.cpu = c;
      s.seg = 0;
      s.tile_begin = 0;
      s.tile_end = 0;
      s.forced_entry = kNone;
      s.entry_hint = kNone;
      // (from the TRUE window, like the tiled geometry: a clamped re-run keeps every slab in place)
      s.rows_cap = (w.head - r.tail) / (sw ? 32 : 40 + 8ull * d.n_events) + 2;
      segs.push_back(s);
      r.n_seg = 1;
    } else {
      const uint64_t t0 = r.tail / kTile;
      // segment geometry is derived from the TRUE head so that a clamped re-run
      // keeps the same segments (and the forced entries keep their meaning)
      const uint64_t t1 = (w.head + kTile - 1) / kTile;
      const uint64_t nt = t1 - t0;
      uint64_t per = (nt + segs_per_cpu - 1) / segs_per_cpu;
      if (per < min_tiles_per_seg) per = min_tiles_per_seg;
      uint32_t k = 0;
      for (uint64_t tb = t0; tb < t1; tb += per, ++k) {
        SegDesc s{};
        s.cpu = c;
        s.seg = k;
        s.tile_begin = tb;
        s.tile_end = tb + per < t1 ? tb + per : t1;
        s.forced_entry = kNone;
        s.entry_hint = kNone;
        if (hint_size && k > 0) {  // first record start at or after the segment's first byte, counted from the tail
          const uint64_t seg_pos = tb * kTile;
          s.entry_hint = r.tail + (seg_pos - r.tail + hint_size - 1) / hint_size * hint_size;
          s.hint_size = hint_size;
        }
        // rows per segment: a SAMPLE is at least 40 + 8n bytes, a handled ContextSwitch record 32
        s.rows_cap = ((s.tile_end - s.tile_begin) * kTile) / (sw ? 32 : 40 + 8ull * d.n_events) + 2;
        segs.push_back(s);
      }
      r.n_seg = k;
    }
    if (with_rows)
      for (uint32_t k = 0; k < r.n_seg; ++k)
        rows_total_words += (1 + segs[r.seg_begin + k].rows_cap) * row_words;
  }
  if (keep_forced && keep_forced->size() == segs.size())
    for (size_t i = 0; i < segs.size(); ++i) segs[i].forced_entry = (*keep_forced)[i].forced_entry;

  if (with_rows) {
    CU(ctx->d_rows.reserve(rows_total_words ? rows_total_words : 1));
    uint64_t off = 0;
    for (uint32_t c = 0; c < C; ++c) {
      RingDesc& r = ctx->h_rings.p[c];
      r.rows = ctx->d_rows.p + off;
      for (uint32_t k = 0; k < r.n_seg; ++k) {
        SegDesc& s = segs[r.seg_begin + k];
        s.rows = ctx->d_rows.p + off;
        off += (1 + s.rows_cap) * row_words;
      }
    }
  }
  ctx->n_segs = (uint32_t)segs.size();
  const size_t ns = segs.size() ? segs.size() : 1;
  CU(ctx->h_segs.reserve(ns));
  CU(ctx->d_segs.reserve(ns));
  CU(ctx->d_results.reserve(ns));
  CU(ctx->h_results.reserve(ns));
  CU(ctx->d_seam.reserve(ns));
  CU(ctx->d_forced.reserve(ns));
  CU(ctx->h_seam.reserve(ns));
  CU(ctx->d_run_list.reserve(ns));
  CU(ctx->d_seg_flag.reserve(ns));
  CU(ctx->h_run_list.reserve(ns));
  if (!segs.empty()) memcpy(ctx->h_segs.p, segs.data(), segs.size() * sizeof(SegDesc));
  return 0;  // k_prepare (run_once) moves the tables to the device
}

// Everything a decode needs set up on the device, in one launch and without the copy
// engines (a context's small table uploads must not queue behind another context's ring
// uploads): tables are read straight from the pinned host staging buffers, scratch words
// are reset, and the histogram cursors are snapshotted for a possible re-run.
struct PrepareParams {
  const uint64_t* src_rings; uint64_t* dst_rings; uint32_t n_ring_words;
  const uint64_t* src_segs;  uint64_t* dst_segs;  uint32_t n_seg_words;
  uint32_t* flags;
  uint64_t* clamp_head; uint32_t n_cpus;
  uint64_t* forced; uint32_t n_forced;
  const uint32_t* cursors; uint32_t* cursors_snap; uint32_t n_lists;  // n_lists = 0: no snapshot
  const uint32_t* src_run; uint32_t* dst_run; uint32_t n_run;
  uint8_t* cpu_done; uint32_t n_done_reset;  // first attempt: every CPU is open again
  uint32_t* seg_flag; uint32_t n_seg_flags;  // k_walk's look-back flags: all of them, or those of the re-run's segments
};

static __global__ void k_prepare(PrepareParams p) {
  const uint32_t stride = gridDim.x * blockDim.x;
  const uint32_t t0 = blockIdx.x * blockDim.x + threadIdx.x;
  for (uint32_t i = t0; i < p.n_seg_words; i += stride) p.dst_segs[i] = p.src_segs[i];
  for (uint32_t i = t0; i < p.n_ring_words; i += stride) p.dst_rings[i] = p.src_rings[i];
  for (uint32_t i = t0; i < p.n_cpus; i += stride) p.clamp_head[i] = ~0ull;
  for (uint32_t i = t0; i < p.n_forced; i += stride) p.forced[i] = ~0ull;
  for (uint32_t i = t0; i < p.n_lists; i += stride) p.cursors_snap[i] = p.cursors[i];
  for (uint32_t i = t0; i < p.n_run; i += stride) p.dst_run[i] = p.src_run[i];
  for (uint32_t i = t0; i < p.n_done_reset; i += stride) p.cpu_done[i] = 0;
  if (p.n_run) {
    for (uint32_t i = t0; i < p.n_run; i += stride) p.seg_flag[p.src_run[i]] = 0;
  } else {
    for (uint32_t i = t0; i < p.n_seg_flags; i += stride) p.seg_flag[i] = 0;
  }
  if (t0 < 4) p.flags[t0] = 0;
}

template <int AGG>
static int launch_walk(hbt_ctx* ctx, const DecodeParams& p, uint32_t n_launch) {
  if constexpr (AGG != HBT_AGG_IP_HISTOGRAM) {
    if (!ctx->walk_attr_set[AGG]) {  // six CTAs' staging slices per SM
      CU(cudaFuncSetAttribute(k_walk<AGG>, cudaFuncAttributePreferredSharedMemoryCarveout,
                              (int)cudaSharedmemCarveoutMaxShared));
      ctx->walk_attr_set[AGG] = true;
    }
    k_walk<AGG><<<(n_launch + kWalkWarps - 1) / kWalkWarps, 32 * kWalkWarps, kWalkWarps * walk_smem_per_warp(AGG),
                  ctx->stream>>>(p, n_launch);
  }
  return 0;
}

template <int AGG>
int launch_decode_agg(hbt_ctx* ctx) {
  DecodeParams p{};
  p.rings = ctx->d_rings.p;
  p.segs = ctx->d_segs.p;
  p.results = ctx->d_results.p;
  p.state_in = ctx->state[ctx->cur].p;
  p.d = ctx->desc;
  p.hist.entries = ctx->hist_entries.p;
  p.hist.cursors = ctx->hist_cursors.p;
  p.hist.buckets = ctx->hist_buckets;
  p.hist.cap = ctx->hist_cap;
  p.hist.overflow = ctx->hist_overflow.p;
  p.cpu_base = ctx->cpu_base;
  p.run_list = ctx->n_run ? ctx->d_run_list.p : nullptr;
  p.seg_flag = ctx->walk ? ctx->d_seg_flag.p : nullptr;
  if (ctx->n_segs == 0) return 0;
  const uint32_t n_launch = ctx->n_run ? ctx->n_run : ctx->n_segs;
  if (ctx->walk) {
    CU(cudaEventRecord(ctx->ev_k0, ctx->stream));
    if (int rc = launch_walk<AGG>(ctx, p, n_launch)) return rc;
    CU(cudaEventRecord(ctx->ev_k1, ctx->stream));
  } else if (ctx->tiled) {
    const size_t smem = decode_smem_bytes(AGG);
    if (!ctx->attr_set[AGG]) {
      CU(cudaFuncSetAttribute(k_decode<AGG>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                              (int)smem));
      // the staging area decides how many CTAs share an SM: ask for all the shared memory
      CU(cudaFuncSetAttribute(k_decode<AGG>, cudaFuncAttributePreferredSharedMemoryCarveout,
                              (int)cudaSharedmemCarveoutMaxShared));
      ctx->attr_set[AGG] = true;
    }
    CU(cudaEventRecord(ctx->ev_k0, ctx->stream));
    k_decode<AGG><<<n_launch, kThreads, smem, ctx->stream>>>(p);
    CU(cudaEventRecord(ctx->ev_k1, ctx->stream));
  } else {
    k_decode_small<AGG><<<(n_launch + 31) / 32, 32, 0, ctx->stream>>>(p, n_launch);
  }
  ++ctx->launches;
  CU(cudaGetLastError());
  return 0;
}

int launch_decode(hbt_ctx* ctx) {
  switch (ctx->desc.aggregate) {
    case HBT_AGG_NONE: return launch_decode_agg<HBT_AGG_NONE>(ctx);
    case HBT_AGG_COUNT_DELTAS: return launch_decode_agg<HBT_AGG_COUNT_DELTAS>(ctx);
    case HBT_AGG_SWITCH_EVENTS: return launch_decode_agg<HBT_AGG_SWITCH_EVENTS>(ctx);
    default: return launch_decode_agg<HBT_AGG_IP_HISTOGRAM>(ctx);
  }
}

int launch_finalize(hbt_ctx* ctx) {
  const uint32_t C = ctx->n_cpus;
  FinalizeParams f{};
  f.rings = ctx->d_rings.p;
  f.segs = ctx->d_segs.p;
  f.forced_out = ctx->d_forced.p;
  f.results = ctx->d_results.p;
  f.state_in = ctx->state[ctx->cur].p;
  f.state_out = ctx->state[ctx->cur ^ 1].p;
  f.out = ctx->d_out.p;
  f.seam_valid = ctx->d_seam.p;
  f.clamp_head = ctx->d_clamp_head.p;
  f.clamp_ret = ctx->d_clamp_ret.p;
  f.locate_from = ctx->d_locate_from.p;
  f.locate_skip = ctx->d_locate_skip.p;
  f.locate_mode = ctx->d_locate_mode.p;
  f.rows_budget = ctx->has_rows_budget ? ctx->d_rows_budget.p : nullptr;
  f.flags = ctx->d_flags.p;
  f.cpu_done = ctx->d_cpu_done.p;
  f.n_cpus = C;
  f.d = ctx->desc;
  f.agg = (int)ctx->desc.aggregate;
  f.seg_limit_tile = ctx->tiled ? kTile : 0;
  k_finalize<<<C, 32, 0, ctx->stream>>>(f);
  ++ctx->launches;
  CU(cudaGetLastError());
  return 0;
}

int hist_fold(hbt_ctx* ctx);

// Launch decode + finalize (+ the histogram fold) for the tables currently uploaded.
int run_once(hbt_ctx* ctx, bool snapshot_cursors) {
  static_assert(sizeof(RingDesc) % 8 == 0 && sizeof(SegDesc) % 8 == 0, "tables move as u64 words");
  PrepareParams pp{};
  void* dp = nullptr;
  CU(cudaHostGetDevicePointer(&dp, ctx->h_rings.p, 0));
  pp.src_rings = static_cast<const uint64_t*>(dp);
  pp.dst_rings = reinterpret_cast<uint64_t*>(ctx->d_rings.p);
  pp.n_ring_words = ctx->n_cpus * (uint32_t)(sizeof(RingDesc) / 8);
  CU(cudaHostGetDevicePointer(&dp, ctx->h_segs.p, 0));
  pp.src_segs = static_cast<const uint64_t*>(dp);
  pp.dst_segs = reinterpret_cast<uint64_t*>(ctx->d_segs.p);
  pp.n_seg_words = ctx->n_segs * (uint32_t)(sizeof(SegDesc) / 8);
  pp.flags = ctx->d_flags.p;
  pp.clamp_head = ctx->d_clamp_head.p;
  pp.n_cpus = ctx->n_cpus;
  pp.forced = ctx->d_forced.p;
  pp.n_forced = ctx->n_segs ? ctx->n_segs : 1;
  if (snapshot_cursors) {
    pp.cursors = ctx->hist_cursors.p;
    pp.cursors_snap = ctx->hist_cursors_snap.p;
    pp.n_lists = (uint32_t)((size_t)ctx->n_cpus * ctx->hist_buckets);
  }
  pp.cpu_done = ctx->d_cpu_done.p;
  pp.n_done_reset = ctx->n_run ? 0 : ctx->n_cpus;  // the first attempt of a decode opens every CPU
  pp.seg_flag = ctx->d_seg_flag.p;
  pp.n_seg_flags = ctx->n_segs;
  if (ctx->n_run) {
    CU(cudaHostGetDevicePointer(&dp, ctx->h_run_list.p, 0));
    pp.src_run = static_cast<const uint32_t*>(dp);
    pp.dst_run = ctx->d_run_list.p;
    pp.n_run = ctx->n_run;
  }
  const uint32_t work = pp.n_seg_words > pp.n_lists ? pp.n_seg_words : pp.n_lists;
  uint32_t grid = (work + 1023) / 1024;
  if (grid < 1) grid = 1;
  if (grid > (uint32_t)ctx->sm_count * 2) grid = (uint32_t)ctx->sm_count * 2;
  k_prepare<<<grid, 256, 0, ctx->stream>>>(pp);
  ++ctx->launches;
  CU(cudaGetLastError());
  int rc = launch_decode(ctx);
  if (rc) return rc;
  rc = launch_finalize(ctx);
  if (rc) return rc;
  if (ctx->desc.aggregate == HBT_AGG_IP_HISTOGRAM) {
    // the fold follows at once; it skips itself if k_finalize asks for a re-run
    ctx->hist_pending_fold = true;
    rc = hist_fold(ctx);
  }
  return rc;
}

size_t hist_lists(const hbt_ctx* ctx) { return (size_t)ctx->n_cpus * ctx->hist_buckets; }

HistLists hist_view(hbt_ctx* ctx) {
  return HistLists{ctx->hist_entries.p, ctx->hist_cursors.p, ctx->hist_buckets, ctx->hist_cap,
                   ctx->hist_overflow.p};
}

// Fold the raw samples the last decode appended into distinct keys (k_hist_aggregate).
int hist_fold(hbt_ctx* ctx) {
  if (!ctx->hist_pending_fold) return 0;
  const size_t smem_full = agg_smem_bytes(kAggSlots / 2);
  if (!ctx->agg_attr_set) {
    CU(cudaFuncSetAttribute(k_hist_aggregate, cudaFuncAttributeMaxDynamicSharedMemorySize,
                            (int)smem_full));
    ctx->agg_attr_set = true;
  }
  const uint32_t n = (uint32_t)hist_lists(ctx);
  // Lists are folded by size class, smallest first: most lists are short and a small
  // staging area lets many CTAs share an SM; the few long ones (a warm key repeats in
  // its list) get the full-size one.  Shared memory = 24 bytes per entry of the class.
  CU(cudaEventRecord(ctx->ev_f0, ctx->stream));
  uint32_t lo = 1;
  for (uint32_t hi = kAggSlots / 8; hi <= kAggSlots / 2; hi *= 2) {
    const size_t smem = agg_smem_bytes(hi);
    uint32_t per_sm = (uint32_t)((220 * 1024) / (smem + 1024));
    if (per_sm > 8) per_sm = 8;
    uint32_t grid = (uint32_t)ctx->sm_count * per_sm * 4;
    if (grid > n) grid = n;
    k_hist_aggregate<<<grid, kAggThreads, smem, ctx->stream>>>(hist_view(ctx), n, lo, hi, ctx->d_flags.p);
    ++ctx->launches;
    lo = hi;
  }
  CU(cudaEventRecord(ctx->ev_f1, ctx->stream));
  ctx->fold_timed = true;
  CU(cudaGetLastError());
  ctx->hist_pending_fold = false;
  return 0;
}

}  // namespace

static int complete_decode(hbt_ctx* ctx);

// ctx_internal.h
int hbt_ctx_device(const hbt_ctx* ctx) { return ctx->device; }
cudaStream_t hbt_ctx_stream(const hbt_ctx* ctx) { return ctx->stream; }
int hbt_ctx_sm_count(const hbt_ctx* ctx) { return ctx->sm_count; }
void hbt_ctx_count_launches(hbt_ctx* ctx, unsigned n) { ctx->launches += n; }
int hbt_ctx_fail(hbt_ctx* ctx, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  if (ctx) ctx->err = buf;
  return code;
}

extern "C" {

int hbt_abi_version(void) { return HBT_B200_ABI_VERSION; }

int hbt_create(hbt_ctx** out, int device, uint32_t max_cpus, uint32_t max_events) {
  if (!out) return -EINVAL;
  *out = nullptr;
  if (max_cpus == 0 || max_cpus > HBT_MAX_CPUS || max_events == 0 || max_events > HBT_MAX_EVENTS)
    return -EINVAL;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || device < 0 || device >= n) {
    cudaGetLastError();
    return -ENODEV;
  }
  if (cudaSetDevice(device) != cudaSuccess) return -ENODEV;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return -ENODEV;
  if (prop.major < 10) return -ENODEV;  // sm_100a code only
  hbt_ctx* ctx = new hbt_ctx();
  ctx->device = device;
  ctx->sm_count = prop.multiProcessorCount;
  ctx->max_cpus = max_cpus;
  ctx->max_events = max_events;
  if (cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking) != cudaSuccess) {
    delete ctx;
    return -EIO;
  }
  ctx->stream = ctx->own_stream;
  cudaEventCreate(&ctx->ev_k0);
  cudaEventCreate(&ctx->ev_k1);
  cudaEventCreate(&ctx->ev_f0);
  cudaEventCreate(&ctx->ev_f1);
  *out = ctx;
  return 0;
}

void hbt_destroy(hbt_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  ctx->ring_store.release();
  ctx->state[0].release();
  ctx->state[1].release();
  ctx->d_rings.release();
  ctx->d_segs.release();
  ctx->d_results.release();
  ctx->d_out.release();
  ctx->d_seam.release();
  ctx->d_forced.release();
  ctx->d_clamp_head.release();
  ctx->d_locate_from.release();
  ctx->d_locate_skip.release();
  ctx->d_locate_mode.release();
  ctx->d_rows_budget.release();
  ctx->d_clamp_ret.release();
  ctx->d_flags.release();
  ctx->d_rows.release();
  ctx->d_rows_packed.release();
  ctx->d_run_list.release();
  ctx->h_run_list.release();
  ctx->d_cpu_done.release();
  ctx->d_seg_flag.release();
  ctx->h_cpu_done.release();
  ctx->h_slabs.release();
  ctx->h_rings.release();
  ctx->h_segs.release();
  ctx->h_results.release();
  ctx->h_out.release();
  ctx->h_seam.release();
  ctx->h_clamp_head.release();
  ctx->h_clamp_ret.release();
  ctx->h_flags.release();
  ctx->sw_state_ts.release();
  ctx->sw_state_err.release();
  ctx->d_sw.release();
  ctx->d_sw_off.release();
  ctx->h_sw_off.release();
  ctx->sw_events.release();
  ctx->sw_merge[0].release();
  ctx->sw_merge[1].release();
  ctx->d_merge_prefix.release();
  ctx->d_merge_cuts.release();
  ctx->h_merge_prefix.release();
  ctx->hist_entries.release();
  ctx->hist_cursors.release();
  ctx->hist_cursors_snap.release();
  ctx->hist_overflow.release();
  ctx->hist_out.release();
  ctx->hist_n.release();
  ctx->hist_list_off.release();
  ctx->hist_cpu_total.release();
  ctx->h_hist_off.release();
  ctx->hist_packed.release();
  comm_release(ctx);
  if (ctx->comm_stream) {
    cudaStreamDestroy(ctx->comm_stream);
    for (int k = 0; k < kMaxCommChunks; ++k) {
      cudaEventDestroy(ctx->ev_chunk_reduced[k]);
      cudaEventDestroy(ctx->ev_chunk_merged[k]);
    }
  }
  ctx->d_out_all.release();
  ctx->h_out_all.release();
  ctx->d_hist_counts.release();
  ctx->h_hist_counts.release();
  ctx->d_ratios.release();
  ctx->c_reads.release();
  ctx->c_sums.release();
  ctx->c_scaled.release();
  ctx->c_deltas.release();
  ctx->c_rates.release();
  ctx->c_ratio.release();
  if (ctx->ev_k0) cudaEventDestroy(ctx->ev_k0);
  if (ctx->ev_k1) cudaEventDestroy(ctx->ev_k1);
  if (ctx->ev_f0) cudaEventDestroy(ctx->ev_f0);
  if (ctx->ev_f1) cudaEventDestroy(ctx->ev_f1);
  if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
  delete ctx;
}

const char* hbt_last_error(const hbt_ctx* ctx) { return ctx ? ctx->err.c_str() : "no context"; }

int hbt_set_stream(hbt_ctx* ctx, void* cuda_stream) {
  if (!ctx) return -EINVAL;
  // a pending decode finishes on the stream it was queued on (its completion reads results
  // back and may launch re-runs there)
  if (int rc = complete_decode(ctx)) return rc;
  ctx->stream = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->own_stream;
  return 0;
}

int hbt_sync(hbt_ctx* ctx) {
  if (!ctx) return -EINVAL;
  CU(cudaStreamSynchronize(ctx->stream));
  return 0;
}

uint64_t hbt_launch_count(const hbt_ctx* ctx) { return ctx ? ctx->launches : 0; }

int hbt_set_geometry(hbt_ctx* ctx, uint32_t segs_per_cpu, uint32_t min_tiles_per_seg) {
  if (!ctx) return -EINVAL;
  if (segs_per_cpu > (uint32_t)kMaxSegsPerCpu) return fail(ctx, -EINVAL, "segs_per_cpu > %d", kMaxSegsPerCpu);
  ctx->geo_segs_per_cpu = segs_per_cpu;
  ctx->geo_min_tiles = min_tiles_per_seg;
  return 0;
}

int hbt_set_decode_kernel(hbt_ctx* ctx, int which) {
  if (!ctx) return -EINVAL;
  if (which != HBT_KERNEL_AUTO && which != HBT_KERNEL_TILES && which != HBT_KERNEL_WALK)
    return fail(ctx, -EINVAL, "unknown decode kernel %d", which);
  ctx->kernel_choice = which;
  return 0;
}

int hbt_set_cpu_base(hbt_ctx* ctx, uint32_t cpu_base) {
  if (!ctx) return -EINVAL;
  ctx->cpu_base = cpu_base;
  return 0;
}

// ---- rings -----------------------------------------------------------------

int hbt_ring_reset_state(hbt_ctx* ctx, uint64_t seed_tstamp) {
  if (!ctx || ctx->n_cpus == 0) return fail(ctx, -EINVAL, "no rings configured");
  if (int rc = complete_decode(ctx)) return rc;
  k_seed_state<<<(ctx->n_cpus + 63) / 64, 64, 0, ctx->stream>>>(ctx->state[ctx->cur].p, ctx->n_cpus,
                                                               seed_tstamp);
  ++ctx->launches;
  CU(cudaGetLastError());
  // CpuThreadSwitchGenerator starts with last_tstamp_ = numeric_limits<TimeStamp>::min() and
  // in_write_errors_ = false (PerCpuThreadSwitchGenerator.h:437-438)
  CU(cudaMemsetAsync(ctx->sw_state_ts.p, 0, ctx->n_cpus * sizeof(uint64_t), ctx->stream));
  CU(cudaMemsetAsync(ctx->sw_state_err.p, 0, ctx->n_cpus * sizeof(uint32_t), ctx->stream));
  for (auto& w : ctx->win) w.active = false;
  ctx->have_results = false;
  ctx->sw_valid = false;
  return 0;
}

int hbt_ring_set_carry(hbt_ctx* ctx, uint32_t cpu, uint64_t carry_tstamp, const uint64_t* last,
                       uint32_t n_events, const uint64_t* counters, uint64_t num_unknown) {
  if (!ctx || cpu >= ctx->n_cpus) return fail(ctx, -EINVAL, "cpu out of range");
  if (n_events > HBT_MAX_EVENTS || (n_events && !last)) return fail(ctx, -EINVAL, "bad last_[]");
  CU(cudaSetDevice(ctx->device));
  if (int rc = complete_decode(ctx)) return rc;
  // the committed generator state of this CPU; the copies below are ordered on the ctx stream
  // behind everything queued so far and read their (pageable or pinned) sources before returning
  uint8_t* st = reinterpret_cast<uint8_t*>(ctx->state[ctx->cur].p + cpu);
  CU(cudaMemcpyAsync(st + offsetof(CpuState, carry_tstamp), &carry_tstamp, 8, cudaMemcpyHostToDevice,
                     ctx->stream));
  CU(cudaMemcpyAsync(st + offsetof(CpuState, num_unknown), &num_unknown, 8, cudaMemcpyHostToDevice,
                     ctx->stream));
  if (n_events)
    CU(cudaMemcpyAsync(st + offsetof(CpuState, carry_last), last, 8ull * n_events,
                       cudaMemcpyHostToDevice, ctx->stream));
  if (counters)
    CU(cudaMemcpyAsync(st + offsetof(CpuState, counters), counters, 8ull * HBT_NUM_COUNTERS,
                       cudaMemcpyHostToDevice, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return 0;
}

int hbt_set_rows_budget(hbt_ctx* ctx, const uint64_t* rows_per_cpu, uint32_t n_cpus) {
  if (!ctx || ctx->n_cpus == 0) return fail(ctx, -EINVAL, "no rings configured");
  if (int rc = complete_decode(ctx)) return rc;
  if (!rows_per_cpu) {
    ctx->has_rows_budget = false;
    return 0;
  }
  if (n_cpus != ctx->n_cpus) return fail(ctx, -EINVAL, "one budget per configured CPU (%u), got %u", ctx->n_cpus, n_cpus);
  CU(cudaSetDevice(ctx->device));
  CU(cudaMemcpyAsync(ctx->d_rows_budget.p, rows_per_cpu, n_cpus * sizeof(uint64_t), cudaMemcpyHostToDevice,
                     ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  ctx->has_rows_budget = true;
  return 0;
}

int hbt_ring_set_switch_state(hbt_ctx* ctx, uint32_t cpu, uint64_t last_tstamp, uint32_t in_write_errors) {
  if (!ctx || cpu >= ctx->n_cpus) return fail(ctx, -EINVAL, "cpu out of range");
  CU(cudaSetDevice(ctx->device));
  if (int rc = complete_decode(ctx)) return rc;
  const uint32_t e = in_write_errors ? 1u : 0u;
  CU(cudaMemcpyAsync(ctx->sw_state_ts.p + cpu, &last_tstamp, 8, cudaMemcpyHostToDevice, ctx->stream));
  CU(cudaMemcpyAsync(ctx->sw_state_err.p + cpu, &e, 4, cudaMemcpyHostToDevice, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return 0;
}

static int configure_common(hbt_ctx* ctx, uint32_t n_cpus) {
  if (int rc = complete_decode(ctx)) return rc;
  if (n_cpus != ctx->hist_n_cpus) {
    // the histogram lists are laid out [n_cpus][lists][cap]: a different CPU count needs
    // hbt_hist_configure again (validate_desc refuses an IP_HISTOGRAM decode until then)
    ctx->hist_buckets = 0;
    ctx->hist_n_cpus = 0;
  }
  ctx->n_cpus = n_cpus;
  ctx->win.assign(n_cpus, Window());
  CU(ctx->state[0].reserve(n_cpus));
  CU(ctx->state[1].reserve(n_cpus));
  CU(ctx->d_out.reserve(n_cpus));
  CU(ctx->h_out.reserve(n_cpus));
  CU(ctx->sw_state_ts.reserve(n_cpus));
  CU(ctx->sw_state_err.reserve(n_cpus));
  CU(ctx->d_sw_off.reserve(n_cpus + 1));
  CU(ctx->h_sw_off.reserve(n_cpus + 1));
  CU(ctx->d_clamp_head.reserve(n_cpus));
  CU(ctx->d_clamp_ret.reserve(n_cpus));
  CU(ctx->d_locate_from.reserve(n_cpus));
  CU(ctx->d_locate_skip.reserve(n_cpus));
  CU(ctx->d_locate_mode.reserve(n_cpus));
  CU(ctx->d_rows_budget.reserve(n_cpus));
  ctx->has_rows_budget = false;
  CU(ctx->h_clamp_head.reserve(n_cpus));
  CU(ctx->h_clamp_ret.reserve(n_cpus));
  CU(ctx->d_flags.reserve(4));
  CU(ctx->h_flags.reserve(8));
  CU(ctx->d_cpu_done.reserve(n_cpus));
  CU(ctx->h_cpu_done.reserve(n_cpus));
  ctx->cur = 0;
  return hbt_ring_reset_state(ctx, kInvalidTs);
}

int hbt_ring_configure(hbt_ctx* ctx, uint32_t n_cpus, uint64_t data_size) {
  if (!ctx) return -EINVAL;
  if (n_cpus == 0 || n_cpus > ctx->max_cpus) return fail(ctx, -EINVAL, "n_cpus %u", n_cpus);
  if (data_size != 0 && (!is_pow2(data_size) || data_size < 4096))
    return fail(ctx, -EINVAL, "data_size must be a power of two >= 4096");
  CU(cudaSetDevice(ctx->device));
  ctx->data_size = data_size;
  if (data_size) {
    CU(ctx->ring_store.reserve((size_t)n_cpus * data_size));
  }
  return configure_common(ctx, n_cpus);
}

int hbt_submit_ring(hbt_ctx* ctx, uint32_t cpu, const void* p1, size_t n1, const void* p2,
                    size_t n2, uint64_t tail) {
  if (!ctx || cpu >= ctx->n_cpus) return fail(ctx, -EINVAL, "cpu out of range");
  if (int rc = complete_decode(ctx)) return rc;  // a possible re-run still reads the image
  if (ctx->data_size == 0 || !ctx->ring_store.p)
    return fail(ctx, -EINVAL, "hbt_ring_configure with a data_size first");
  const uint64_t ds = ctx->data_size, mask = ds - 1;
  if ((tail & 7) || (n1 & 7) || (n2 & 7)) return fail(ctx, -EINVAL, "unaligned slice");
  if ((uint64_t)n1 + n2 > ds) return fail(ctx, -EINVAL, "slices larger than the ring");
  uint8_t* img = ctx->ring_store.p + (size_t)cpu * ds;
  const uint64_t off = tail & mask;
  if (off + n1 > ds) return fail(ctx, -EINVAL, "first slice crosses the ring end");
  if (n1) CU(cudaMemcpyAsync(img + off, p1, n1, cudaMemcpyHostToDevice, ctx->stream));
  if (n2) {
    // the second slice starts at data_start (RbDataSlices, CpuEventsGroup.h:1559-1561)
    if (off + n1 != ds) return fail(ctx, -EINVAL, "second slice without a wrap");
    if (n2 > off) return fail(ctx, -EINVAL, "second slice overlaps the first");
    CU(cudaMemcpyAsync(img, p2, n2, cudaMemcpyHostToDevice, ctx->stream));
  }
  Window& w = ctx->win[cpu];
  w.base = img;
  w.data_size = ds;
  w.tail = tail;
  w.head = tail + n1 + n2;
  w.active = true;
  return 0;
}

int hbt_submit_ring_dev(hbt_ctx* ctx, uint32_t cpu, const void* ring_base, uint64_t data_size,
                        uint64_t tail, uint64_t head) {
  if (!ctx || cpu >= ctx->n_cpus) return fail(ctx, -EINVAL, "cpu out of range");
  if (int rc = complete_decode(ctx)) return rc;
  if (!ring_base || ((uintptr_t)ring_base & 15)) return fail(ctx, -EINVAL, "ring_base alignment");
  if (!is_pow2(data_size) || data_size < 4096) return fail(ctx, -EINVAL, "bad data_size");
  if ((tail & 7) || head < tail) return fail(ctx, -EINVAL, "bad window");
  Window& w = ctx->win[cpu];
  w.base = static_cast<const uint8_t*>(ring_base);
  w.data_size = data_size;
  w.tail = tail;
  w.head = head;
  w.active = true;
  return 0;
}

int hbt_submit_rings_dev(hbt_ctx* ctx, uint32_t n, const void* const* ring_bases,
                         uint64_t data_size, const uint64_t* tails, const uint64_t* heads) {
  if (!ctx || n > ctx->n_cpus || !ring_bases || !tails || !heads) return fail(ctx, -EINVAL, "bad args");
  if (int rc = complete_decode(ctx)) return rc;
  if (!is_pow2(data_size) || data_size < 4096) return fail(ctx, -EINVAL, "bad data_size");
  for (uint32_t c = 0; c < n; ++c) {
    if (!ring_bases[c] || ((uintptr_t)ring_bases[c] & 15) || (tails[c] & 7) || heads[c] < tails[c])
      return fail(ctx, -EINVAL, "bad ring %u", c);
  }
  for (uint32_t c = 0; c < n; ++c) {
    Window& w = ctx->win[c];
    w.base = static_cast<const uint8_t*>(ring_bases[c]);
    w.data_size = data_size;
    w.tail = tails[c];
    w.head = heads[c];
    w.active = true;
  }
  return 0;
}

// ---- decode ------------------------------------------------------------------

int hbt_decode(hbt_ctx* ctx, const hbt_decode_desc* desc) {
  if (!ctx || ctx->n_cpus == 0) return fail(ctx, -EINVAL, "no rings configured");
  int rc = validate_desc(ctx, desc);
  if (rc) return rc;
  CU(cudaSetDevice(ctx->device));
  rc = complete_decode(ctx);  // an unfetched decode still commits its state
  if (rc) return rc;
  ctx->desc = *desc;
  if (ctx->desc.aggregate != HBT_AGG_COUNT_DELTAS) ctx->desc.n_events = 0;
  if (ctx->desc.aggregate == HBT_AGG_IP_HISTOGRAM) {
    // 32-bit counts in the lists: refuse anything that could wrap one
    uint64_t bound = 0;
    for (const Window& w : ctx->win)
      if (w.active && w.head > w.tail) bound += (w.head - w.tail) / ctx->desc.ip_sample_bytes + 1;
    if (ctx->hist_samples + bound > 0xFFFFFFFFull)
      return fail(ctx, -EOVERFLOW, "more than 2^32-1 samples since hbt_hist_clear");
    ctx->hist_samples += bound;
  }
  rc = build_tables(ctx, nullptr, nullptr, nullptr);
  if (rc) return rc;
  ctx->n_run = 0;
  ctx->rerun_segments = 0;
  // histogram: k_prepare also snapshots the lists' cursors, a re-run starts again from there
  rc = run_once(ctx, ctx->desc.aggregate == HBT_AGG_IP_HISTOGRAM);
  if (rc) return rc;
  ctx->pending = true;
  ctx->have_results = false;
  ctx->rows_packed_valid = false;
  ctx->reruns = 0;
  return 0;
}

}  // extern "C"

// Complete the pending decode: read the re-run flag; while set, take the failed
// attempt back out of the histogram, apply forced entries / clamps and run again.
static int complete_decode(hbt_ctx* ctx) {
  if (!ctx->pending) return 0;
  const uint32_t C = ctx->n_cpus;
  std::vector<uint64_t> clamp_head(C, kNone);
  std::vector<int64_t> clamp_ret(C, 0);
  bool have_clamp = false;
  // Results travel with the re-run flag (one synchronisation in the common case); they only
  // count when the flag is clear.  The event-stream aggregate still has its second pass to run
  // on them, so it copies them afterwards.
  const bool early_results = ctx->desc.aggregate != HBT_AGG_SWITCH_EVENTS;
  auto copy_results = [&]() -> int {
    ctx->h_flags.p[4] = 0;
    if (ctx->desc.aggregate == HBT_AGG_IP_HISTOGRAM)
      CU(cudaMemcpyAsync(ctx->h_flags.p + 4, ctx->hist_overflow.p, sizeof(uint32_t),
                         cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaMemcpyAsync(ctx->h_out.p, ctx->d_out.p, C * sizeof(hbt_cpu_result),
                       cudaMemcpyDeviceToHost, ctx->stream));
    if (ctx->n_segs) {
      CU(cudaMemcpyAsync(ctx->h_results.p, ctx->d_results.p, ctx->n_segs * sizeof(SegResult),
                         cudaMemcpyDeviceToHost, ctx->stream));
      CU(cudaMemcpyAsync(ctx->h_seam.p, ctx->d_seam.p, ctx->n_segs, cudaMemcpyDeviceToHost,
                         ctx->stream));
    }
    return 0;
  };
  for (int iter = 0;; ++iter) {
    CU(cudaMemcpyAsync(ctx->h_flags.p, ctx->d_flags.p, 4 * sizeof(uint32_t),
                       cudaMemcpyDeviceToHost, ctx->stream));
    if (early_results)
      if (int rc = copy_results()) return rc;
    CU(cudaStreamSynchronize(ctx->stream));
    if (ctx->h_flags.p[0] == 0) break;
    if (iter > 4 * kMaxSegsPerCpu + 8) return fail(ctx, -EIO, "decode re-run did not converge");
    ++ctx->reruns;
    // locate cuts for a finite max_num_records, then fetch clamps, forced entries and which CPUs are settled
    k_locate<<<(C + 63) / 64, 64, 0, ctx->stream>>>(ctx->d_rings.p, C, ctx->d_clamp_head.p,
                                                   ctx->d_locate_from.p, ctx->d_locate_skip.p,
                                                   ctx->d_locate_mode.p, ctx->desc.handled_mask);
    ++ctx->launches;
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(ctx->h_clamp_head.p, ctx->d_clamp_head.p, C * sizeof(uint64_t),
                       cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaMemcpyAsync(ctx->h_clamp_ret.p, ctx->d_clamp_ret.p, C * sizeof(int64_t),
                       cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaMemcpyAsync(ctx->h_cpu_done.p, ctx->d_cpu_done.p, C, cudaMemcpyDeviceToHost, ctx->stream));
    std::vector<SegDesc> forced(ctx->h_segs.p, ctx->h_segs.p + ctx->n_segs);
    std::vector<uint64_t> forced_new(ctx->n_segs, kNone);
    if (ctx->n_segs)
      CU(cudaMemcpyAsync(forced_new.data(), ctx->d_forced.p, ctx->n_segs * sizeof(uint64_t),
                         cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    // Only what the failed attempt got wrong is decoded again.  A CPU whose chain held is settled
    // (its results stay).  Of an open CPU: a clamped window (max_num_records, a regression at a
    // seam, an error under the histogram) changes where every segment stops -> all its segments;
    // the histogram cannot take single segments back out of the shared lists -> all its segments,
    // after its cursors are put back; otherwise only the segment whose entry guess was wrong --
    // the other segments' results stand and are chained again.
    const bool hist = ctx->desc.aggregate == HBT_AGG_IP_HISTOGRAM;
    uint32_t n_run = 0;
    for (uint32_t c = 0; c < C; ++c) {
      if (ctx->h_cpu_done.p[c]) continue;
      const RingDesc& r = ctx->h_rings.p[c];
      bool whole = hist;
      if (ctx->h_clamp_head.p[c] != kNone) {
        clamp_head[c] = ctx->h_clamp_head.p[c];
        clamp_ret[c] = ctx->h_clamp_ret.p[c];
        have_clamp = true;
        whole = true;
      }
      for (uint32_t k = 0; k < r.n_seg; ++k) {
        const uint32_t gi = r.seg_begin + k;
        if (forced_new[gi] != kNone) forced[gi].forced_entry = forced_new[gi];
        if (whole || forced_new[gi] != kNone) ctx->h_run_list.p[n_run++] = gi;
      }
      if (hist && r.n_seg) {
        // forget what the failed attempt appended to this CPU's lists
        const size_t l0 = (size_t)c * ctx->hist_buckets;
        CU(cudaMemcpyAsync(ctx->hist_cursors.p + l0, ctx->hist_cursors_snap.p + l0,
                           ctx->hist_buckets * sizeof(uint32_t), cudaMemcpyDeviceToDevice, ctx->stream));
      }
    }
    if (n_run == 0) return fail(ctx, -EIO, "decode asked for a re-run without naming a segment");
    int rc = build_tables(ctx, have_clamp ? clamp_head.data() : nullptr, clamp_ret.data(), &forced);
    if (rc) return rc;
    ctx->n_run = n_run;
    ctx->rerun_segments += n_run;
    rc = run_once(ctx, false);
    if (rc) return rc;
  }
  // success: fold the appended samples, commit the generator state, fetch results
  // (histogram: the fold of the successful run was queued by run_once)
  ctx->cur ^= 1;
  ctx->sw_valid = false;
  ctx->sw_merged = nullptr;
  SwParams sp{};
  if (ctx->desc.aggregate == HBT_AGG_SWITCH_EVENTS) {
    // pass 2 of the event stream: counts and offsets now, the events once their number is known
    CU(ctx->d_sw.reserve(ctx->n_segs ? ctx->n_segs : 1));
    sp.rings = ctx->d_rings.p;
    sp.segs = ctx->d_segs.p;
    sp.results = ctx->d_results.p;
    sp.seg_flags = ctx->d_seam.p;
    sp.sw = ctx->d_sw.p;
    sp.out = ctx->d_out.p;
    sp.state_ts = ctx->sw_state_ts.p;
    sp.state_err = ctx->sw_state_err.p;
    sp.offsets = ctx->d_sw_off.p;
    sp.n_cpus = C;
    sp.n_segs = ctx->n_segs;
    sp.cpu_base = ctx->cpu_base;
    k_sw_scan<<<1, 1024, 0, ctx->stream>>>(sp);
    ++ctx->launches;
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(ctx->h_sw_off.p, ctx->d_sw_off.p, (C + 1) * sizeof(uint64_t),
                       cudaMemcpyDeviceToHost, ctx->stream));
  }
  if (!early_results) {
    if (int rc = copy_results()) return rc;
    CU(cudaStreamSynchronize(ctx->stream));
  }
  // consumed windows: tail advances like writeDataTail (CpuEventsGroup.h:1520)
  for (uint32_t c = 0; c < C; ++c) {
    Window& w = ctx->win[c];
    if (w.active) w.tail = ctx->h_out.p[c].new_tail;
  }
  if (ctx->desc.aggregate == HBT_AGG_SWITCH_EVENTS) {
    const uint64_t total = ctx->h_sw_off.p[C];
    CU(ctx->sw_events.reserve(total ? total : 1));
    sp.events = ctx->sw_events.p;
    if (ctx->n_segs && total) {
      k_sw_emit<<<ctx->n_segs, 256, 0, ctx->stream>>>(sp);
      ++ctx->launches;
      CU(cudaGetLastError());
    }
    ctx->sw_valid = true;
  }
  ctx->pending = false;
  ctx->have_results = true;
  return 0;
}

extern "C" {

int hbt_fetch_results(hbt_ctx* ctx, hbt_cpu_result* out, uint32_t cap) {
  if (!ctx || !out) return -EINVAL;
  int rc = complete_decode(ctx);
  if (rc) return rc;
  if (!ctx->have_results) return fail(ctx, -EINVAL, "no decode to fetch");
  if (cap < ctx->n_cpus) return fail(ctx, -ENOSPC, "cap %u < n_cpus %u", cap, ctx->n_cpus);
  memcpy(out, ctx->h_out.p, ctx->n_cpus * sizeof(hbt_cpu_result));
  // the overflow flag travelled with the results (everything stays on the context's stream:
  // a copy on the legacy stream would queue behind other contexts' uploads)
  if (ctx->desc.aggregate == HBT_AGG_IP_HISTOGRAM && ctx->h_flags.p[4])
    return fail(ctx, -ENOSPC, "a histogram list is full: hbt_hist_configure more capacity");
  return (int)ctx->n_cpus;
}

uint32_t hbt_decode_reruns(const hbt_ctx* ctx) { return ctx ? ctx->reruns : 0; }
uint64_t hbt_decode_rerun_segments(const hbt_ctx* ctx) { return ctx ? ctx->rerun_segments : 0; }

float hbt_decode_kernel_ms(hbt_ctx* ctx) {
  if (!ctx || !ctx->tiled) return -1.0f;
  float ms = -1.0f;
  if (cudaEventSynchronize(ctx->ev_k1) != cudaSuccess) return -1.0f;
  if (cudaEventElapsedTime(&ms, ctx->ev_k0, ctx->ev_k1) != cudaSuccess) {
    cudaGetLastError();
    return -1.0f;
  }
  return ms;
}

float hbt_hist_fold_kernel_ms(hbt_ctx* ctx) {
  if (!ctx || !ctx->fold_timed) return -1.0f;
  float ms = -1.0f;
  if (cudaEventSynchronize(ctx->ev_f1) != cudaSuccess) return -1.0f;
  if (cudaEventElapsedTime(&ms, ctx->ev_f0, ctx->ev_f1) != cudaSuccess) {
    cudaGetLastError();
    return -1.0f;
  }
  return ms;
}

uint64_t hbt_decode_slow_tiles(hbt_ctx* ctx) {
  if (!ctx || !ctx->have_results) return 0;
  uint64_t n = 0;
  for (uint32_t i = 0; i < ctx->n_segs; ++i) n += ctx->h_results.p[i].slow_tiles;
  return n;
}

uint32_t hbt_rows_max_slabs(const hbt_ctx* ctx) { (void)ctx; return 2 * kMaxSegsPerCpu; }

int hbt_rows_slabs_dev(hbt_ctx* ctx, uint32_t cpu, const uint64_t** ptr, uint64_t* n_rows,
                       uint32_t* n_slabs) {
  if (!ctx || cpu >= ctx->n_cpus || !ptr || !n_rows || !n_slabs) return -EINVAL;
  int rc = complete_decode(ctx);
  if (rc) return rc;
  if (!ctx->have_results || ctx->desc.aggregate != HBT_AGG_COUNT_DELTAS)
    return fail(ctx, -EINVAL, "no COUNT_DELTAS decode to read");
  const RingDesc& r = ctx->h_rings.p[cpu];
  const uint64_t row_words = 2 + (uint64_t)ctx->desc.n_events;
  uint32_t k = 0;
  uint64_t remaining = ctx->h_out.p[cpu].n_rows;
  for (uint32_t s = 0; s < r.n_seg && remaining; ++s) {
    const uint32_t gi = r.seg_begin + s;
    const SegDesc& sd = ctx->h_segs.p[gi];
    if (ctx->h_seam.p[gi] & 1) {
      ptr[k] = sd.rows;
      n_rows[k] = 1;
      ++k;
      --remaining;
    }
    uint64_t n = ctx->h_results.p[gi].n_rows;
    if (n > remaining) n = remaining;
    if (n) {
      ptr[k] = sd.rows + row_words;
      n_rows[k] = n;
      ++k;
      remaining -= n;
    }
  }
  *n_slabs = k;
  return 0;
}

int64_t hbt_fetch_rows(hbt_ctx* ctx, uint32_t cpu, uint64_t* rows, uint64_t cap_rows) {
  if (!ctx || cpu >= ctx->n_cpus) return -EINVAL;
  const uint64_t* ptr[2 * kMaxSegsPerCpu];
  uint64_t n[2 * kMaxSegsPerCpu];
  uint32_t k = 0;
  int rc = hbt_rows_slabs_dev(ctx, cpu, ptr, n, &k);
  if (rc) return rc;
  const uint64_t total = ctx->h_out.p[cpu].n_rows;
  if (total > cap_rows) return fail(ctx, -ENOSPC, "cap_rows %llu < %llu",
                                    (unsigned long long)cap_rows, (unsigned long long)total);
  const uint64_t row_words = 2 + (uint64_t)ctx->desc.n_events;
  uint64_t off = 0;
  for (uint32_t i = 0; i < k; ++i) {
    CU(cudaMemcpyAsync(rows + off * row_words, ptr[i], n[i] * row_words * 8,
                       cudaMemcpyDeviceToHost, ctx->stream));
    off += n[i];
  }
  CU(cudaStreamSynchronize(ctx->stream));
  return (int64_t)off;
}

}  // extern "C"

// Packs the row slabs of a decode (one or two runs per segment: the seam row, the segment's
// rows) into one array, CPU after CPU in stream order: one CTA per run, consecutive threads copy
// consecutive u64 words.  The run table lives in pinned host memory (read over PCIe once).
static __global__ void __launch_bounds__(256)
    k_rows_pack(const RowSlab* __restrict__ slabs, uint32_t n_slabs, uint32_t row_words,
                uint64_t* __restrict__ dst) {
  for (uint32_t s = blockIdx.x; s < n_slabs; s += gridDim.x) {
    const RowSlab sl = slabs[s];
    const uint64_t n = sl.n_rows * row_words;
    uint64_t* out = dst + sl.dst_row * row_words;
    for (uint64_t w = threadIdx.x; w < n; w += blockDim.x) out[w] = sl.src[w];
  }
}

extern "C" {

int hbt_rows_packed_dev(hbt_ctx* ctx, const uint64_t** rows, uint64_t* offsets) {
  if (!ctx || !rows || !offsets) return -EINVAL;
  int rc = complete_decode(ctx);
  if (rc) return rc;
  if (!ctx->have_results || ctx->desc.aggregate != HBT_AGG_COUNT_DELTAS)
    return fail(ctx, -EINVAL, "no COUNT_DELTAS decode to read");
  const uint32_t C = ctx->n_cpus;
  const uint32_t row_words = 2 + ctx->desc.n_events;
  if (!ctx->rows_packed_valid) {
    CU(cudaSetDevice(ctx->device));
    CU(ctx->h_slabs.reserve((size_t)2 * ctx->n_segs + 1));
    ctx->rows_offsets.assign(C + 1, 0);
    uint32_t k = 0;
    uint64_t total = 0;
    for (uint32_t c = 0; c < C; ++c) {  // the same walk as hbt_rows_slabs_dev, for every CPU
      ctx->rows_offsets[c] = total;
      const RingDesc& r = ctx->h_rings.p[c];
      uint64_t remaining = ctx->h_out.p[c].n_rows;
      for (uint32_t sgm = 0; sgm < r.n_seg && remaining; ++sgm) {
        const uint32_t gi = r.seg_begin + sgm;
        const SegDesc& sd = ctx->h_segs.p[gi];
        if (ctx->h_seam.p[gi] & 1) {
          ctx->h_slabs.p[k++] = RowSlab{sd.rows, total, 1};
          ++total;
          --remaining;
        }
        uint64_t n = ctx->h_results.p[gi].n_rows;
        if (n > remaining) n = remaining;
        if (n) {
          ctx->h_slabs.p[k++] = RowSlab{sd.rows + row_words, total, n};
          total += n;
          remaining -= n;
        }
      }
    }
    ctx->rows_offsets[C] = total;
    CU(ctx->d_rows_packed.reserve(total ? total * row_words : 1));
    if (k) {
      void* dp = nullptr;
      CU(cudaHostGetDevicePointer(&dp, ctx->h_slabs.p, 0));
      uint32_t grid = k < (uint32_t)ctx->sm_count * 16 ? k : (uint32_t)ctx->sm_count * 16;
      k_rows_pack<<<grid, 256, 0, ctx->stream>>>(static_cast<const RowSlab*>(dp), k, row_words,
                                                 ctx->d_rows_packed.p);
      ++ctx->launches;
      CU(cudaGetLastError());
    }
    ctx->rows_packed_valid = true;
  }
  *rows = ctx->d_rows_packed.p;
  memcpy(offsets, ctx->rows_offsets.data(), (C + 1) * sizeof(uint64_t));
  return 0;
}

int64_t hbt_fetch_rows_all(hbt_ctx* ctx, uint64_t* rows, uint64_t cap_rows, uint64_t* offsets) {
  if (!ctx || !rows || !offsets) return -EINVAL;
  const uint64_t* dev = nullptr;
  if (int rc = hbt_rows_packed_dev(ctx, &dev, offsets)) return rc;
  const uint64_t total = offsets[ctx->n_cpus];
  if (total > cap_rows)
    return fail(ctx, -ENOSPC, "cap_rows %llu < %llu", (unsigned long long)cap_rows, (unsigned long long)total);
  if (total)
    CU(cudaMemcpyAsync(rows, dev, total * (2 + (uint64_t)ctx->desc.n_events) * 8, cudaMemcpyDeviceToHost,
                       ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));  // also orders the pack kernel's reads of the pinned table
  return (int64_t)total;
}

int64_t hbt_fetch_switch_events(hbt_ctx* ctx, uint32_t cpu, hbt_switch_event* out, uint64_t cap) {
  if (!ctx || cpu >= ctx->n_cpus || !out) return -EINVAL;
  if (int rc = complete_decode(ctx)) return rc;
  if (!ctx->sw_valid) return fail(ctx, -EINVAL, "no SWITCH_EVENTS decode to read");
  const uint64_t b = ctx->h_sw_off.p[cpu], n = ctx->h_sw_off.p[cpu + 1] - b;
  if (n > cap) return fail(ctx, -ENOSPC, "cap %llu < %llu events", (unsigned long long)cap,
                           (unsigned long long)n);
  if (n) CU(cudaMemcpyAsync(out, ctx->sw_events.p + b, n * sizeof(hbt_switch_event),
                            cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return (int64_t)n;
}

int hbt_switch_events_dev(hbt_ctx* ctx, const hbt_switch_event** events, uint64_t* offsets) {
  if (!ctx || !events || !offsets) return -EINVAL;
  if (int rc = complete_decode(ctx)) return rc;
  if (!ctx->sw_valid) return fail(ctx, -EINVAL, "no SWITCH_EVENTS decode to read");
  *events = ctx->sw_events.p;
  memcpy(offsets, ctx->h_sw_off.p, (ctx->n_cpus + 1) * sizeof(uint64_t));
  return 0;
}

int hbt_switch_events_merged_dev(hbt_ctx* ctx, const hbt_switch_event** merged, uint64_t* n_events) {
  if (!ctx || !merged || !n_events) return -EINVAL;
  if (int rc = complete_decode(ctx)) return rc;
  if (!ctx->sw_valid) return fail(ctx, -EINVAL, "no SWITCH_EVENTS decode to read");
  const uint32_t C = ctx->n_cpus;
  const uint64_t total = ctx->h_sw_off.p[C];
  *n_events = total;
  if (!ctx->sw_merged) {
    const hbt_switch_event* cur = ctx->sw_events.p;
    if (total) {
      int which = 0;
      // per level: first CTA of every pair of runs (one CTA per kMergeTile outputs)
      std::vector<uint32_t> prefix;
      std::vector<uint32_t> level_at;
      for (uint32_t run = 1; run < C; run *= 2) {
        level_at.push_back((uint32_t)prefix.size());
        const uint32_t n_pairs = (C + 2 * run - 1) / (2 * run);
        uint32_t acc = 0;
        for (uint32_t p = 0; p < n_pairs; ++p) {
          prefix.push_back(acc);
          const uint32_t lo = std::min(2 * p * run, C), hi = std::min((2 * p + 2) * run, C);
          const uint64_t len = ctx->h_sw_off.p[hi] - ctx->h_sw_off.p[lo];
          acc += (uint32_t)((len + kMergeTile - 1) / kMergeTile);
        }
        prefix.push_back(acc);
      }
      CU(ctx->d_merge_prefix.reserve(prefix.size() ? prefix.size() : 1));
      CU(ctx->h_merge_prefix.reserve(prefix.size() ? prefix.size() : 1));
      memcpy(ctx->h_merge_prefix.p, prefix.data(), prefix.size() * sizeof(uint32_t));
      CU(cudaMemcpyAsync(ctx->d_merge_prefix.p, ctx->h_merge_prefix.p, prefix.size() * sizeof(uint32_t),
                         cudaMemcpyHostToDevice, ctx->stream));
      {
        uint32_t max_grid = 1;
        for (size_t l = 0; l < level_at.size(); ++l) {
          const size_t end = l + 1 < level_at.size() ? level_at[l + 1] : prefix.size();
          max_grid = std::max(max_grid, prefix[end - 1]);
        }
        CU(ctx->d_merge_cuts.reserve(max_grid));
      }
      uint32_t level = 0;
      for (uint32_t run = 1; run < C; run *= 2, which ^= 1, ++level) {
        CU(ctx->sw_merge[which].reserve(total));
        const uint32_t n_pairs = (C + 2 * run - 1) / (2 * run);
        const uint32_t* pre = ctx->d_merge_prefix.p + level_at[level];
        const uint32_t grid = prefix[level_at[level] + n_pairs];
        if (grid) {
          k_sw_cuts<<<(grid + 255) / 256, 256, 0, ctx->stream>>>(reinterpret_cast<const ulonglong2*>(cur),
                                                                 ctx->d_sw_off.p, C, run, pre, n_pairs, grid,
                                                                 ctx->d_merge_cuts.p);
          k_sw_merge<<<grid, 256, 0, ctx->stream>>>(
              reinterpret_cast<const ulonglong2*>(cur), reinterpret_cast<ulonglong2*>(ctx->sw_merge[which].p),
              ctx->d_sw_off.p, C, run, pre, n_pairs, ctx->d_merge_cuts.p);
          ctx->launches += 2;
        }
        cur = ctx->sw_merge[which].p;
      }
      CU(cudaGetLastError());
    }
    ctx->sw_merged = cur;
  }
  *merged = ctx->sw_merged;
  return 0;
}

int64_t hbt_fetch_switch_events_merged(hbt_ctx* ctx, hbt_switch_event* out, uint64_t cap) {
  if (!out) return -EINVAL;
  const hbt_switch_event* m = nullptr;
  uint64_t n = 0;
  if (int rc = hbt_switch_events_merged_dev(ctx, &m, &n)) return rc;
  if (n > cap) return fail(ctx, -ENOSPC, "cap %llu < %llu events", (unsigned long long)cap,
                           (unsigned long long)n);
  if (n) CU(cudaMemcpyAsync(out, m, n * sizeof(hbt_switch_event), cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return (int64_t)n;
}

int hbt_fetch_carry(hbt_ctx* ctx, uint32_t cpu, uint64_t* last, uint32_t n_events) {
  if (!ctx || cpu >= ctx->n_cpus || !last || n_events > HBT_MAX_EVENTS) return -EINVAL;
  int rc = complete_decode(ctx);
  if (rc) return rc;
  CU(cudaMemcpyAsync(last, ctx->state[ctx->cur].p[cpu].carry_last, n_events * 8,
                     cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return 0;
}

// ---- histogram -----------------------------------------------------------------

int hbt_hist_clear(hbt_ctx* ctx) {
  if (!ctx || ctx->hist_buckets == 0) return fail(ctx, -EINVAL, "no histogram configured");
  if (int rc = complete_decode(ctx)) return rc;
  // clearing = emptying the lists: only the cursors are touched
  CU(cudaMemsetAsync(ctx->hist_cursors.p, 0, hist_lists(ctx) * sizeof(uint32_t), ctx->stream));
  CU(cudaMemsetAsync(ctx->hist_overflow.p, 0, sizeof(uint32_t), ctx->stream));
  ctx->hist_samples = 0;
  ctx->hist_pending_fold = false;
  return 0;
}

int hbt_hist_configure(hbt_ctx* ctx, uint64_t capacity_slots) {
  if (!ctx || ctx->n_cpus == 0) return fail(ctx, -EINVAL, "configure rings first");
  if (int rc = complete_decode(ctx)) return rc;
  const uint64_t per_cpu = capacity_slots / ctx->n_cpus;
  const uint32_t cap = kAggSlots / 2;  // a list always fits the fold's shared-memory table
  const uint64_t buckets = (per_cpu + cap - 1) / cap;
  if (per_cpu == 0 || buckets > (1u << 20))
    return fail(ctx, -EINVAL, "capacity %llu gives %llu lists per CPU",
                (unsigned long long)capacity_slots, (unsigned long long)buckets);
  CU(cudaSetDevice(ctx->device));
  ctx->hist_buckets = (uint32_t)buckets;
  ctx->hist_cap = cap;
  ctx->hist_n_cpus = ctx->n_cpus;
  CU(ctx->hist_entries.reserve(hist_lists(ctx) * cap));
  CU(ctx->hist_cursors.reserve(hist_lists(ctx)));
  CU(ctx->hist_cursors_snap.reserve(hist_lists(ctx)));
  CU(ctx->hist_overflow.reserve(1));
  CU(ctx->hist_n.reserve(1));
  return hbt_hist_clear(ctx);
}

static int64_t hist_collect(hbt_ctx* ctx, hbt_hist_entry* out, uint64_t cap, uint32_t first_cpu = 0,
                            uint32_t n_cpus = 0xFFFFFFFFu) {
  if (!ctx || ctx->hist_buckets == 0) return fail(ctx, -EINVAL, "no histogram configured");
  int rc = complete_decode(ctx);
  if (rc) return rc;
  if (n_cpus == 0xFFFFFFFFu) n_cpus = ctx->n_cpus - (first_cpu < ctx->n_cpus ? first_cpu : ctx->n_cpus);
  if (first_cpu + (uint64_t)n_cpus > ctx->n_cpus) return fail(ctx, -EINVAL, "CPU range outside the configured CPUs");
  if (out) CU(ctx->hist_out.reserve(cap ? cap : 1));
  CU(cudaMemsetAsync(ctx->hist_n.p, 0, sizeof(unsigned long long), ctx->stream));
  // the lists of CPUs [first_cpu, first_cpu + n_cpus) are contiguous
  HistLists v = hist_view(ctx);
  v.entries += (uint64_t)first_cpu * ctx->hist_buckets * ctx->hist_cap;
  v.cursors += (uint64_t)first_cpu * ctx->hist_buckets;
  const uint32_t n_lists = n_cpus * ctx->hist_buckets;
  unsigned long long n = 0;
  if (n_lists) {
    uint32_t grid = (uint32_t)ctx->sm_count * 16;
    if (grid > n_lists) grid = n_lists;
    k_hist_gather<<<grid, 128, 0, ctx->stream>>>(v, n_lists, ctx->cpu_base + first_cpu,
                                                 out ? ctx->hist_out.p : nullptr, ctx->hist_n.p, cap);
    ++ctx->launches;
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(&n, ctx->hist_n.p, sizeof(n), cudaMemcpyDeviceToHost, ctx->stream));
  }
  CU(cudaStreamSynchronize(ctx->stream));
  if (out) {
    if (n > cap) return fail(ctx, -ENOSPC, "cap %llu < %llu entries", (unsigned long long)cap, n);
    CU(cudaMemcpyAsync(out, ctx->hist_out.p, n * sizeof(hbt_hist_entry), cudaMemcpyDeviceToHost,
                       ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
  }
  return (int64_t)n;
}

int64_t hbt_hist_size(hbt_ctx* ctx) { return hist_collect(ctx, nullptr, 0); }

int64_t hbt_hist_fetch(hbt_ctx* ctx, hbt_hist_entry* out, uint64_t cap) {
  if (!out) return -EINVAL;
  return hist_collect(ctx, out, cap);
}

int64_t hbt_hist_fetch_cpus(hbt_ctx* ctx, uint32_t first_cpu, uint32_t n_cpus, hbt_hist_entry* out,
                            uint64_t cap) {
  return hist_collect(ctx, out, out ? cap : 0, first_cpu, n_cpus);
}

int64_t hbt_hist_fetch_packed(hbt_ctx* ctx, uint32_t first_cpu, uint32_t n_cpus, hbt_hist_key_count* out,
                              uint64_t cap, uint64_t* offsets) {
  if (!ctx || ctx->hist_buckets == 0) return fail(ctx, -EINVAL, "no histogram configured");
  if (!offsets) return fail(ctx, -EINVAL, "offsets is NULL");
  if (int rc = complete_decode(ctx)) return rc;
  if (n_cpus == 0 || first_cpu + (uint64_t)n_cpus > ctx->n_cpus)
    return fail(ctx, -EINVAL, "CPU range outside the configured CPUs");
  static_assert(sizeof(hbt_hist_key_count) == sizeof(ulonglong2), "the device's list entry IS the wire entry");
  HistLists v = hist_view(ctx);
  v.entries += (uint64_t)first_cpu * ctx->hist_buckets * ctx->hist_cap;
  v.cursors += (uint64_t)first_cpu * ctx->hist_buckets;
  const uint32_t n_lists = n_cpus * ctx->hist_buckets;
  CU(ctx->hist_list_off.reserve(n_lists));
  CU(ctx->hist_cpu_total.reserve(2 * (size_t)HBT_MAX_CPUS + 2));
  CU(ctx->h_hist_off.reserve((size_t)HBT_MAX_CPUS + 1));
  unsigned long long* d_tot = ctx->hist_cpu_total.p;
  unsigned long long* d_off = d_tot + HBT_MAX_CPUS;
  k_hist_offsets<<<n_cpus, 256, 0, ctx->stream>>>(v, ctx->hist_list_off.p, d_tot);
  k_hist_cpu_scan<<<1, 512, 0, ctx->stream>>>(d_tot, n_cpus, d_off);
  ctx->launches += 2;
  CU(cudaGetLastError());
  if (out) {
    // packed on the device before the host knows how many entries there are (cap bounds the writes)
    CU(ctx->hist_packed.reserve(cap ? cap : 1));
    uint32_t grid = (uint32_t)ctx->sm_count * 16;
    if (grid > n_lists) grid = n_lists;
    k_hist_pack<<<grid, 128, 0, ctx->stream>>>(v, n_lists, ctx->hist_list_off.p, d_off, ctx->hist_packed.p, cap);
    ++ctx->launches;
    CU(cudaGetLastError());
  }
  CU(cudaMemcpyAsync(ctx->h_hist_off.p, d_off, (n_cpus + 1) * sizeof(unsigned long long), cudaMemcpyDeviceToHost,
                     ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  for (uint32_t i = 0; i <= n_cpus; ++i) offsets[i] = ctx->h_hist_off.p[i];
  const uint64_t n = offsets[n_cpus];
  if (out) {
    if (n > cap) return fail(ctx, -ENOSPC, "cap %llu < %llu entries", (unsigned long long)cap, (unsigned long long)n);
    CU(cudaMemcpyAsync(out, ctx->hist_packed.p, n * sizeof(hbt_hist_key_count), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
  }
  return (int64_t)n;
}

int hbt_hist_lists_dev(hbt_ctx* ctx, const void** entries, const uint32_t** cursors,
                       uint32_t* lists_per_cpu, uint32_t* list_capacity) {
  if (!ctx || !entries || !cursors || !lists_per_cpu || !list_capacity || ctx->hist_buckets == 0)
    return -EINVAL;
  if (int rc = complete_decode(ctx)) return rc;
  *entries = ctx->hist_entries.p;
  *cursors = ctx->hist_cursors.p;
  *lists_per_cpu = ctx->hist_buckets;
  *list_capacity = ctx->hist_cap;
  return 0;
}

// ---- counter groups -----------------------------------------------------------------

static int counts_check(hbt_ctx* ctx, uint64_t n_intervals, uint32_t n_cpus, uint32_t n_events,
                        uint32_t n_ratios, const hbt_ratio* ratios) {
  if (!ctx) return -EINVAL;
  if (n_cpus == 0 || n_cpus > HBT_MAX_CPUS) return fail(ctx, -EINVAL, "n_cpus %u", n_cpus);
  if (n_events == 0 || n_events > ctx->max_events) return fail(ctx, -EINVAL, "n_events %u", n_events);
  if (n_ratios > kMaxRatios) return fail(ctx, -EINVAL, "n_ratios %u > %d", n_ratios, kMaxRatios);
  for (uint32_t r = 0; r < n_ratios; ++r)
    if (!ratios || ratios[r].num_event >= n_events || ratios[r].den_event >= n_events)
      return fail(ctx, -EINVAL, "ratio %u names an event out of range", r);
  (void)n_intervals;
  return 0;
}

// ---- metric descriptors ---------------------------------------------------------------

uint32_t hbt_metric_count(void) { return kNumMetrics; }

const hbt_metric_desc* hbt_metric_at(uint32_t index) {
  return index < kNumMetrics ? &kMetricTable[index] : nullptr;
}

const hbt_metric_desc* hbt_metric_find(const char* id, const char* arch) {
  if (!id) return nullptr;
  const hbt_metric_desc* any = nullptr;
  for (uint32_t i = 0; i < kNumMetrics; ++i) {
    const hbt_metric_desc& d = kMetricTable[i];
    if (strcmp(d.id, id) != 0) continue;
    if (arch && strcmp(d.arch, arch) == 0) return &d;   // architecture-specific list first
    if (strcmp(d.arch, "any") == 0) any = &d;
  }
  return any;
}

int hbt_metric_event_index(const hbt_metric_desc* desc, const char* nickname) {
  if (!desc || !nickname) return -EINVAL;
  for (uint32_t i = 0; i < desc->n_events; ++i)
    if (strcmp(desc->nicknames[i], nickname) == 0) return (int)i;
  return -ENOENT;
}

static int upload_ratios(hbt_ctx* ctx, const hbt_ratio* ratios, uint32_t n_ratios) {
  CU(ctx->d_ratios.reserve(kMaxRatios));
  if (n_ratios)
    CU(cudaMemcpyAsync(ctx->d_ratios.p, ratios, n_ratios * sizeof(hbt_ratio),
                       cudaMemcpyHostToDevice, ctx->stream));
  return 0;
}

static int counts_grid(const hbt_ctx* ctx, uint64_t n_intervals) {
  // one warp per interval, 8 warps per CTA, 8 CTAs per SM
  uint64_t g = (uint64_t)ctx->sm_count * 8;
  const uint64_t need = (n_intervals + 7) / 8;
  if (g > need) g = need;
  return (int)(g ? g : 1);
}

static int counts_launch(hbt_ctx* ctx, const uint64_t* reads, uint64_t n_intervals,
                         uint32_t n_cpus, uint32_t n_events, uint32_t n_ratios, uint64_t* sums,
                         uint64_t* scaled, float* rates, double* ratio_out, uint64_t* deltas,
                         int do_scale, const hbt_ratio* dev_ratios = nullptr) {
  if (n_intervals == 0) return 0;
  CountsArgs a{};
  a.reads = reads;
  a.n_intervals = n_intervals;
  a.n_cpus = n_cpus;
  a.n_events = n_events;
  a.ratios = dev_ratios ? dev_ratios : ctx->d_ratios.p;
  a.n_ratios = n_ratios;
  a.sums = sums;
  a.rates = rates;
  a.ratio_out = ratio_out;
  a.do_scale = do_scale;
  a.scaled = scaled;
  if (do_scale && deltas && !scaled) {
    CU(ctx->c_scaled.reserve(n_intervals * n_events));
    a.scaled = ctx->c_scaled.p;
  }
  const uint64_t interval_bytes = (uint64_t)n_cpus * (3 + n_events) * 8;
  const bool tma_ok = 3 + n_events <= 32 && interval_bytes <= (uint64_t)kCountsStageBytes &&
                      (interval_bytes & 15) == 0 && ((uintptr_t)reads & 15) == 0;
  if (tma_ok) {
    // whole intervals staged in shared memory by bulk async copies
    const size_t smem = (size_t)kCountsStages * kCountsStageBytes + 64 + 8 * 32 * 8;
    if (!ctx->counts_attr_set) {
      CU(cudaFuncSetAttribute(k_counts_reduce_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      ctx->counts_attr_set = true;
    }
    const uint32_t per_stage = (uint32_t)(kCountsStageBytes / interval_bytes);
    uint64_t grid = (uint64_t)ctx->sm_count * 2;
    const uint64_t chunks = (n_intervals + per_stage - 1) / per_stage;
    if (grid > chunks) grid = chunks;
    k_counts_reduce_tma<<<(unsigned)grid, 256, smem, ctx->stream>>>(a, per_stage);
  } else if (3 + n_events <= 32) {
    k_counts_reduce<false><<<counts_grid(ctx, n_intervals), 256, 0, ctx->stream>>>(a);
  } else {
    k_counts_reduce<true><<<counts_grid(ctx, n_intervals), 256, 0, ctx->stream>>>(a);
  }
  ++ctx->launches;
  CU(cudaGetLastError());
  if (do_scale && deltas) {
    const uint64_t total = n_intervals * n_events;
    int g = (int)((total + 255) / 256);
    if (g > ctx->sm_count * 16) g = ctx->sm_count * 16;
    k_scaled_diff<<<g, 256, 0, ctx->stream>>>(a.scaled, nullptr, n_intervals, n_events, deltas);
    ++ctx->launches;
    CU(cudaGetLastError());
  }
  return 0;
}

int hbt_counts_reduce_dev(hbt_ctx* ctx, const uint64_t* reads, uint64_t n_intervals,
                          uint32_t n_cpus, uint32_t n_events, uint64_t* sums) {
  int rc = counts_check(ctx, n_intervals, n_cpus, n_events, 0, nullptr);
  if (rc) return rc;
  if (!reads || !sums) return fail(ctx, -EINVAL, "NULL buffer");
  CU(cudaSetDevice(ctx->device));
  return counts_launch(ctx, reads, n_intervals, n_cpus, n_events, 0, sums, nullptr, nullptr,
                       nullptr, nullptr, 0);
}

int hbt_counts_scale_dev(hbt_ctx* ctx, const uint64_t* sums, uint64_t n_intervals,
                         uint32_t n_events, const hbt_ratio* ratios, uint32_t n_ratios,
                         const uint64_t* prev_scaled, uint64_t* scaled, float* rates,
                         double* ratio_out, uint64_t* deltas) {
  int rc = counts_check(ctx, n_intervals, 1, n_events, n_ratios, ratios);
  if (rc) return rc;
  if (!sums) return fail(ctx, -EINVAL, "NULL buffer");
  if (n_intervals == 0) return 0;
  CU(cudaSetDevice(ctx->device));
  rc = upload_ratios(ctx, ratios, n_ratios);
  if (rc) return rc;
  ScaleArgs a{};
  a.sums = sums;
  a.n_intervals = n_intervals;
  a.n_events = n_events;
  a.ratios = ctx->d_ratios.p;
  a.n_ratios = ratio_out ? n_ratios : 0;
  a.scaled = scaled;
  if (deltas && !scaled) {
    CU(ctx->c_scaled.reserve(n_intervals * n_events));
    a.scaled = ctx->c_scaled.p;
  }
  a.rates = rates;
  a.ratio_out = ratio_out;
  const uint64_t total = n_intervals * (n_events + a.n_ratios);
  int g = (int)((total + 255) / 256);
  if (g > ctx->sm_count * 16) g = ctx->sm_count * 16;
  k_counts_scale<<<g, 256, 0, ctx->stream>>>(a);
  ++ctx->launches;
  CU(cudaGetLastError());
  if (deltas) {
    k_scaled_diff<<<g, 256, 0, ctx->stream>>>(a.scaled, prev_scaled, n_intervals, n_events, deltas);
    ++ctx->launches;
    CU(cudaGetLastError());
  }
  return 0;
}

int hbt_counts_pipeline_dev(hbt_ctx* ctx, const uint64_t* reads, uint64_t n_intervals,
                            uint32_t n_cpus, uint32_t n_events, const hbt_ratio* ratios,
                            uint32_t n_ratios, uint64_t* sums, uint64_t* scaled, float* rates,
                            double* ratio_out, uint64_t* deltas) {
  int rc = counts_check(ctx, n_intervals, n_cpus, n_events, n_ratios, ratios);
  if (rc) return rc;
  if (!reads) return fail(ctx, -EINVAL, "NULL buffer");
  CU(cudaSetDevice(ctx->device));
  rc = upload_ratios(ctx, ratios, n_ratios);
  if (rc) return rc;
  return counts_launch(ctx, reads, n_intervals, n_cpus, n_events, ratio_out ? n_ratios : 0, sums,
                       scaled, rates, ratio_out, deltas, 1);
}

int hbt_counts_pipeline_metric_dev(hbt_ctx* ctx, const hbt_metric_desc* desc, const uint64_t* reads,
                                   uint64_t n_intervals, uint32_t n_cpus, uint64_t* sums,
                                   uint64_t* scaled, float* rates, double* ratio_out,
                                   uint64_t* deltas) {
  if (!ctx) return -EINVAL;
  if (!desc || desc < kMetricTable || desc >= kMetricTable + kNumMetrics)
    return fail(ctx, -EINVAL, "desc is not an entry of the library's metric table");
  int rc = counts_check(ctx, n_intervals, n_cpus, desc->n_events, 0, nullptr);
  if (rc) return rc;
  if (!reads) return fail(ctx, -EINVAL, "NULL buffer");
  CU(cudaSetDevice(ctx->device));
  if (!ctx->metric_ratios_dev) {
    void* sym = nullptr;  // device address of the constant table on this ctx's device
    CU(cudaGetSymbolAddress(&sym, c_metric_ratios));
    ctx->metric_ratios_dev = static_cast<const hbt_ratio*>(sym);
  }
  return counts_launch(ctx, reads, n_intervals, n_cpus, desc->n_events,
                       ratio_out ? desc->n_ratios : 0, sums, scaled, rates, ratio_out, deltas, 1,
                       ctx->metric_ratios_dev + (size_t)desc->index * kMaxMetricRatios);
}

int hbt_counts_pipeline_host(hbt_ctx* ctx, const uint64_t* reads, uint64_t n_intervals,
                             uint32_t n_cpus, uint32_t n_events, const hbt_ratio* ratios,
                             uint32_t n_ratios, uint64_t* sums, uint64_t* scaled, float* rates,
                             double* ratio_out, uint64_t* deltas) {
  int rc = counts_check(ctx, n_intervals, n_cpus, n_events, n_ratios, ratios);
  if (rc) return rc;
  if (!reads) return fail(ctx, -EINVAL, "NULL buffer");
  if (n_intervals == 0) return 0;
  CU(cudaSetDevice(ctx->device));
  const uint64_t W = 3 + (uint64_t)n_events;
  const uint64_t n_in = n_intervals * n_cpus * W;
  CU(ctx->c_reads.reserve(n_in));
  CU(ctx->c_sums.reserve(n_intervals * W));
  CU(ctx->c_scaled.reserve(n_intervals * n_events));
  CU(ctx->c_deltas.reserve(n_intervals * n_events));
  CU(ctx->c_rates.reserve(n_intervals * n_events));
  CU(ctx->c_ratio.reserve(n_intervals * (n_ratios ? n_ratios : 1)));
  CU(cudaMemcpyAsync(ctx->c_reads.p, reads, n_in * 8, cudaMemcpyHostToDevice, ctx->stream));
  rc = upload_ratios(ctx, ratios, n_ratios);
  if (rc) return rc;
  rc = counts_launch(ctx, ctx->c_reads.p, n_intervals, n_cpus, n_events,
                     ratio_out ? n_ratios : 0, sums ? ctx->c_sums.p : nullptr, ctx->c_scaled.p,
                     rates ? ctx->c_rates.p : nullptr, ratio_out ? ctx->c_ratio.p : nullptr,
                     deltas ? ctx->c_deltas.p : nullptr, 1);
  if (rc) return rc;
  if (sums)
    CU(cudaMemcpyAsync(sums, ctx->c_sums.p, n_intervals * W * 8, cudaMemcpyDeviceToHost, ctx->stream));
  if (scaled)
    CU(cudaMemcpyAsync(scaled, ctx->c_scaled.p, n_intervals * n_events * 8, cudaMemcpyDeviceToHost,
                       ctx->stream));
  if (rates)
    CU(cudaMemcpyAsync(rates, ctx->c_rates.p, n_intervals * n_events * 4, cudaMemcpyDeviceToHost,
                       ctx->stream));
  if (ratio_out && n_ratios)
    CU(cudaMemcpyAsync(ratio_out, ctx->c_ratio.p, n_intervals * n_ratios * 8,
                       cudaMemcpyDeviceToHost, ctx->stream));
  if (deltas)
    CU(cudaMemcpyAsync(deltas, ctx->c_deltas.p, n_intervals * n_events * 8, cudaMemcpyDeviceToHost,
                       ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return 0;
}

}  // extern "C"

// ---- multi-GPU merge over NCCL ---------------------------------------------
//
// One process per GPU; the ranks of a box shard the CPUs.  The decode path needs no exchange
// (per-CPU outputs are disjoint by key); what IS exchanged: (i) the partial group sums of
// counter groups -- an integer SUM all-reduce BEFORE the scaling (PerCpuBase.h:104-119 sums,
// CpuEventsGroup.h:449-459 scales the sum) -- and (ii) gathers of per-CPU-keyed outputs.
// libnccl is bound at run time (dlopen; the copy the process already holds, e.g. torch's, is
// preferred), so the library has no link-time dependency on it and a single-GPU daemon never
// loads it.

namespace {

struct NcclUniqueId {  // ncclUniqueId: passed by value
  char b[HBT_COMM_ID_BYTES];
};

struct NcclApi {
  // prototypes of nccl.h (NCCL 2.x ABI): results are ints (0 = ncclSuccess), comms opaque
  int (*GetUniqueId)(void* id) = nullptr;
  int (*CommInitRank)(void** comm, int nranks, NcclUniqueId id, int rank) = nullptr;
  int (*CommDestroy)(void* comm) = nullptr;
  int (*AllReduce)(const void* send, void* recv, size_t count, int dtype, int op, void* comm,
                   cudaStream_t stream) = nullptr;
  int (*AllGather)(const void* send, void* recv, size_t sendcount, int dtype, void* comm,
                   cudaStream_t stream) = nullptr;
  int (*Send)(const void* send, size_t count, int dtype, int peer, void* comm, cudaStream_t stream) = nullptr;
  int (*Recv)(void* recv, size_t count, int dtype, int peer, void* comm, cudaStream_t stream) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  bool ok = false;
};
constexpr int kNcclSum = 0, kNcclUint8 = 1, kNcclUint64 = 5;  // ncclRedOp_t / ncclDataType_t values

NcclApi* nccl_api() {
  static NcclApi api;
  static bool tried = false;
  if (tried) return api.ok ? &api : nullptr;
  tried = true;
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);  // the copy already in the process
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) return nullptr;
  auto sym = [&](const char* n) { return dlsym(h, n); };
  api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
  api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
  api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
  api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(sym("ncclAllReduce"));
  api.AllGather = reinterpret_cast<decltype(api.AllGather)>(sym("ncclAllGather"));
  api.Send = reinterpret_cast<decltype(api.Send)>(sym("ncclSend"));
  api.Recv = reinterpret_cast<decltype(api.Recv)>(sym("ncclRecv"));
  api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(sym("ncclGroupStart"));
  api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(sym("ncclGroupEnd"));
  api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
  api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.AllReduce && api.AllGather &&
           api.Send && api.Recv && api.GroupStart && api.GroupEnd && api.GetErrorString;
  return api.ok ? &api : nullptr;
}

int nccl_fail(hbt_ctx* ctx, int rc, const char* what) {
  NcclApi* a = nccl_api();
  return fail(ctx, -EIO, "%s: %s", what, a ? a->GetErrorString(rc) : "libnccl.so.2 not loadable");
}
#define NC(call)                                      \
  do {                                                \
    int r__ = (call);                                 \
    if (r__ != 0) return nccl_fail(ctx, r__, #call);  \
  } while (0)

// word0 of a merged group sum: n_events if every rank's shard read fine (each contributed
// n_events), else 0 -- a failed read anywhere fails the interval (PerCpuBase.h:111-113)
__global__ void k_fix_word0(uint64_t* sums, uint64_t n_intervals, uint32_t W, uint64_t n_events,
                            uint64_t n_ranks) {
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n_intervals;
       i += (uint64_t)gridDim.x * blockDim.x)
    sums[i * W] = sums[i * W] == n_ranks * n_events ? n_events : 0;
}

void comm_release(hbt_ctx* ctx) {
  if (ctx->comm && ctx->comm_owned)
    if (NcclApi* a = nccl_api()) a->CommDestroy(ctx->comm);
  ctx->comm = nullptr;
}

int comm_ready(hbt_ctx* ctx) {
  if (!ctx) return -EINVAL;
  if (!ctx->comm) return fail(ctx, -ENOTCONN, "no communicator: hbt_comm_init or hbt_comm_attach first");
  return 0;
}

int comm_streams(hbt_ctx* ctx) {
  if (!ctx->comm_stream) {
    CU(cudaStreamCreateWithFlags(&ctx->comm_stream, cudaStreamNonBlocking));
    for (int k = 0; k < kMaxCommChunks; ++k) {
      CU(cudaEventCreateWithFlags(&ctx->ev_chunk_reduced[k], cudaEventDisableTiming));
      CU(cudaEventCreateWithFlags(&ctx->ev_chunk_merged[k], cudaEventDisableTiming));
    }
  }
  return 0;
}

int allreduce_sums(hbt_ctx* ctx, uint64_t* sums, uint64_t n_intervals, uint32_t n_events, cudaStream_t st) {
  NcclApi* a = nccl_api();
  const uint32_t W = 3 + n_events;
  NC(a->AllReduce(sums, sums, n_intervals * W, kNcclUint64, kNcclSum, ctx->comm, st));
  int g = (int)((n_intervals + 255) / 256);
  if (g > ctx->sm_count * 8) g = ctx->sm_count * 8;
  k_fix_word0<<<g, 256, 0, st>>>(sums, n_intervals, W, n_events, (uint64_t)ctx->n_ranks);
  ++ctx->launches;
  CU(cudaGetLastError());
  return 0;
}

}  // namespace

extern "C" {

int hbt_comm_unique_id(void* id_out) {
  NcclApi* a = nccl_api();
  if (!a || !id_out) return a ? -EINVAL : -ENOSYS;
  return a->GetUniqueId(id_out) == 0 ? 0 : -EIO;
}

int hbt_comm_init(hbt_ctx* ctx, const void* id, int n_ranks, int rank) {
  if (!ctx || !id || n_ranks < 1 || rank < 0 || rank >= n_ranks) return fail(ctx, -EINVAL, "bad rank / n_ranks");
  NcclApi* a = nccl_api();
  if (!a) return fail(ctx, -ENOSYS, "libnccl.so.2 could not be loaded");
  if (ctx->comm) return fail(ctx, -EALREADY, "this context already has a communicator");
  CU(cudaSetDevice(ctx->device));
  NcclUniqueId uid;
  memcpy(uid.b, id, HBT_COMM_ID_BYTES);
  NC(a->CommInitRank(&ctx->comm, n_ranks, uid, rank));
  ctx->comm_owned = true;
  ctx->n_ranks = n_ranks;
  ctx->rank = rank;
  return 0;
}

int hbt_comm_attach(hbt_ctx* ctx, void* nccl_comm, int n_ranks, int rank) {
  if (!ctx || !nccl_comm || n_ranks < 1 || rank < 0 || rank >= n_ranks) return fail(ctx, -EINVAL, "bad arguments");
  if (!nccl_api()) return fail(ctx, -ENOSYS, "libnccl.so.2 could not be loaded");
  if (ctx->comm) return fail(ctx, -EALREADY, "this context already has a communicator");
  ctx->comm = nccl_comm;
  ctx->comm_owned = false;
  ctx->n_ranks = n_ranks;
  ctx->rank = rank;
  return 0;
}

int hbt_comm_n_ranks(const hbt_ctx* ctx) { return ctx && ctx->comm ? ctx->n_ranks : 0; }

int hbt_counts_allreduce_dev(hbt_ctx* ctx, uint64_t* sums, uint64_t n_intervals, uint32_t n_events) {
  if (int rc = comm_ready(ctx)) return rc;
  if (!sums || n_events == 0 || n_events > HBT_MAX_EVENTS) return fail(ctx, -EINVAL, "bad arguments");
  if (n_intervals == 0) return 0;
  CU(cudaSetDevice(ctx->device));
  return allreduce_sums(ctx, sums, n_intervals, n_events, ctx->stream);
}

int hbt_counts_pipeline_sharded_dev(hbt_ctx* ctx, const uint64_t* reads, uint64_t n_intervals,
                                    uint32_t n_cpus_local, uint32_t n_events, const hbt_ratio* ratios,
                                    uint32_t n_ratios, uint64_t* sums, uint64_t* scaled, float* rates,
                                    double* ratio_out, uint64_t* deltas) {
  if (int rc = comm_ready(ctx)) return rc;
  int rc = counts_check(ctx, n_intervals, n_cpus_local, n_events, n_ratios, ratios);
  if (rc) return rc;
  if (!reads || !sums) return fail(ctx, -EINVAL, "NULL buffer (sums is needed: it is what the ranks exchange)");
  if (n_intervals == 0) return 0;
  CU(cudaSetDevice(ctx->device));
  if ((rc = comm_streams(ctx))) return rc;
  if ((rc = upload_ratios(ctx, ratios, n_ratios))) return rc;
  const uint64_t W = 3 + (uint64_t)n_events;
  if (deltas && !scaled) {
    CU(ctx->c_scaled.reserve(n_intervals * n_events));
    scaled = ctx->c_scaled.p;
  }
  // Interval chunks: the collective of chunk k (own stream) runs while chunk k+1 is reduced.
  // Chunks of >= 2 MiB of sums per rank keep the all-reduce out of its latency-bound regime (an
  // all-reduce over 8 ranks costs ~100 us whatever its size below a few MiB: eight 4-MiB chunks made the
  // exchange half of config 5's step at N=8).
  uint64_t n_chunks = n_intervals * W * 8 / ((uint64_t)(2u << 20) * (uint64_t)std::max(ctx->n_ranks, 2));
  if (n_chunks < 1) n_chunks = 1;
  if (n_chunks > (uint64_t)kMaxCommChunks) n_chunks = kMaxCommChunks;
  const uint64_t per = (n_intervals + n_chunks - 1) / n_chunks;
  auto scale_chunk = [&](uint64_t k) -> int {
    const uint64_t i0 = k * per, n = std::min(per, n_intervals - i0);
    CU(cudaStreamWaitEvent(ctx->stream, ctx->ev_chunk_merged[k], 0));
    ScaleArgs a{};
    a.sums = sums + i0 * W;
    a.n_intervals = n;
    a.n_events = n_events;
    a.ratios = ctx->d_ratios.p;
    a.n_ratios = ratio_out ? n_ratios : 0;
    a.scaled = scaled ? scaled + i0 * n_events : nullptr;
    a.rates = rates ? rates + i0 * n_events : nullptr;
    a.ratio_out = ratio_out ? ratio_out + i0 * n_ratios : nullptr;
    const uint64_t total = n * (n_events + a.n_ratios);
    int g = (int)((total + 255) / 256);
    if (g > ctx->sm_count * 16) g = ctx->sm_count * 16;
    k_counts_scale<<<g, 256, 0, ctx->stream>>>(a);
    ++ctx->launches;
    CU(cudaGetLastError());
    if (deltas) {
      // interval i0's delta reaches back into the previous chunk, scaled one step earlier on this stream
      k_scaled_diff<<<g, 256, 0, ctx->stream>>>(a.scaled, i0 ? a.scaled - n_events : nullptr, n, n_events,
                                                deltas + i0 * n_events);
      ++ctx->launches;
      CU(cudaGetLastError());
    }
    return 0;
  };
  for (uint64_t k = 0; k < n_chunks; ++k) {
    const uint64_t i0 = k * per;
    if (i0 >= n_intervals) { n_chunks = k; break; }
    const uint64_t n = std::min(per, n_intervals - i0);
    rc = counts_launch(ctx, reads + i0 * n_cpus_local * W, n, n_cpus_local, n_events, 0, sums + i0 * W,
                       nullptr, nullptr, nullptr, nullptr, 0);
    if (rc) return rc;
    CU(cudaEventRecord(ctx->ev_chunk_reduced[k], ctx->stream));
    CU(cudaStreamWaitEvent(ctx->comm_stream, ctx->ev_chunk_reduced[k], 0));
    if ((rc = allreduce_sums(ctx, sums + i0 * W, n, n_events, ctx->comm_stream))) return rc;
    CU(cudaEventRecord(ctx->ev_chunk_merged[k], ctx->comm_stream));
    if (k > 0 && (rc = scale_chunk(k - 1))) return rc;  // behind reduce k: overlaps all-reduce k
  }
  return scale_chunk(n_chunks - 1);
}

int hbt_fetch_results_all(hbt_ctx* ctx, hbt_cpu_result* out_all, uint32_t cap) {
  if (int rc = comm_ready(ctx)) return rc;
  if (!out_all) return -EINVAL;
  if (int rc = complete_decode(ctx)) return rc;
  if (!ctx->have_results) return fail(ctx, -EINVAL, "no decode to fetch");
  const uint32_t C = ctx->n_cpus, total = C * (uint32_t)ctx->n_ranks;
  if (cap < total) return fail(ctx, -ENOSPC, "cap %u < %u results", cap, total);
  CU(cudaSetDevice(ctx->device));
  CU(ctx->d_out_all.reserve(total));
  CU(ctx->h_out_all.reserve(total));
  NcclApi* a = nccl_api();
  // every rank's per-CPU results (disjoint by CPU id) -> all ranks; ranks hold equal CPU counts
  NC(a->AllGather(ctx->d_out.p, ctx->d_out_all.p, C * sizeof(hbt_cpu_result), kNcclUint8, ctx->comm, ctx->stream));
  CU(cudaMemcpyAsync(ctx->h_out_all.p, ctx->d_out_all.p, total * sizeof(hbt_cpu_result), cudaMemcpyDeviceToHost,
                     ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  memcpy(out_all, ctx->h_out_all.p, total * sizeof(hbt_cpu_result));
  if (ctx->desc.aggregate == HBT_AGG_IP_HISTOGRAM && ctx->h_flags.p[4])
    return fail(ctx, -ENOSPC, "a histogram list is full: hbt_hist_configure more capacity");
  return (int)total;
}

int64_t hbt_hist_gather(hbt_ctx* ctx, int root, hbt_hist_entry* out, uint64_t cap, uint64_t* counts_per_rank) {
  if (int rc = comm_ready(ctx)) return rc;
  if (root < 0 || root >= ctx->n_ranks) return fail(ctx, -EINVAL, "root %d", root);
  if (ctx->hist_buckets == 0) return fail(ctx, -EINVAL, "no histogram configured");
  if (int rc = complete_decode(ctx)) return rc;
  CU(cudaSetDevice(ctx->device));
  NcclApi* a = nccl_api();
  const int R = ctx->n_ranks;
  // local entries into one array (device), their number to everybody
  CU(ctx->d_hist_counts.reserve(R + 1));
  CU(ctx->h_hist_counts.reserve(R + 1));
  const uint32_t n_lists = (uint32_t)hist_lists(ctx);
  CU(cudaMemsetAsync(ctx->hist_n.p, 0, sizeof(unsigned long long), ctx->stream));
  uint32_t grid = (uint32_t)ctx->sm_count * 16;
  if (grid > n_lists) grid = n_lists;
  k_hist_gather<<<grid, 128, 0, ctx->stream>>>(hist_view(ctx), n_lists, ctx->cpu_base, nullptr, ctx->hist_n.p, 0);
  ++ctx->launches;
  CU(cudaGetLastError());
  NC(a->AllGather(ctx->hist_n.p, ctx->d_hist_counts.p, 1, kNcclUint64, ctx->comm, ctx->stream));
  CU(cudaMemcpyAsync(ctx->h_hist_counts.p, ctx->d_hist_counts.p, R * sizeof(unsigned long long),
                     cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  uint64_t total = 0, mine_off = 0;
  for (int r = 0; r < R; ++r) {
    if (r == ctx->rank) mine_off = total;
    if (counts_per_rank) counts_per_rank[r] = ctx->h_hist_counts.p[r];
    total += ctx->h_hist_counts.p[r];
  }
  const uint64_t mine = ctx->h_hist_counts.p[ctx->rank];
  const bool is_root = ctx->rank == root;
  if (is_root && (!out || total > cap)) return fail(ctx, -ENOSPC, "cap %llu < %llu entries", (unsigned long long)cap,
                                                   (unsigned long long)total);
  CU(ctx->hist_out.reserve(is_root ? (total ? total : 1) : (mine ? mine : 1)));
  CU(cudaMemsetAsync(ctx->hist_n.p, 0, sizeof(unsigned long long), ctx->stream));
  k_hist_gather<<<grid, 128, 0, ctx->stream>>>(hist_view(ctx), n_lists, ctx->cpu_base,
                                               ctx->hist_out.p + (is_root ? mine_off : 0), ctx->hist_n.p, mine);
  ++ctx->launches;
  CU(cudaGetLastError());
  NC(a->GroupStart());
  if (is_root) {
    uint64_t off = 0;
    for (int r = 0; r < R; ++r) {
      const uint64_t n = ctx->h_hist_counts.p[r];
      if (r != root && n)
        NC(a->Recv(ctx->hist_out.p + off, n * sizeof(hbt_hist_entry), kNcclUint8, r, ctx->comm, ctx->stream));
      off += n;
    }
  } else if (mine) {
    NC(a->Send(ctx->hist_out.p, mine * sizeof(hbt_hist_entry), kNcclUint8, root, ctx->comm, ctx->stream));
  }
  NC(a->GroupEnd());
  if (is_root && total)
    CU(cudaMemcpyAsync(out, ctx->hist_out.p, total * sizeof(hbt_hist_entry), cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return is_root ? (int64_t)total : (int64_t)mine;
}

}  // extern "C"

