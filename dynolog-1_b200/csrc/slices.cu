// Slice aggregation on the device: the interval join that cuts a compute unit's
// slice stream at the count intervals, the per-stack-id slice frequencies, and the interval x bin
// duration matrix.
//
// Replaces, for slices and intervals kept in HBM:
//   IntervalSlicer::makeIntervalIntersectingSlices / consumeSlices   hbt/src/tagstack/IntervalSlicer.cpp:52-216
//   CuMonData::stackIdFreqs, SliceFreq::accum, MonData::stackIdFreqs  hbt/src/mon/MonData.h:30-48, 243-268, 386-403
//   binmap::IntervalBinMatrix, BinMapper, TagStackIdBinner            hbt/src/mon/MonData.h:425-612
//
// The join.  The reference walks the slice ring once per interval, from a consumer mark that
// advances by what the previous interval consumed.  For intervals in increasing time order over a
// time-ordered slice stream the mark never passes the first slice an interval needs, so every
// interval's answer depends on the slices alone: the first slice whose end is not before the
// interval's start (binary search), the slices that start before its end (binary search), the
// abort rules on what lies between, and a trimmed copy of the two boundary slices.  One warp per
// interval; the per-interval slice lists are laid out back to back by a prefix sum.
//
// Frequencies.  An open-addressing table keyed by stack id (64-bit CAS claims a slot) accumulates
// duration and num_obs with u64 atomic adds; "number of distinct intervals a stack id appears in"
// is counted by a second table of (slot, interval) pairs: the thread whose CAS inserts a pair is
// the one occurrence that counts.  All sums are integer, so the result does not depend on order.
#include <cuda_runtime.h>

#include <algorithm>
#include <cerrno>
#include <cstdint>
#include <cstring>
#include <vector>

#include "../../include/hbt_b200.h"
#include "ctx_internal.h"

namespace {

constexpr uint64_t kEmptyKey = ~0ull;  // also tagstack::slice::kInvalidTagStackId: kept in an extra slot
constexpr uint8_t kAnalysis = 1;       // Slice::TransitionType::Analysis (Slicer.h:31-43)

static_assert(sizeof(hbt_slice) == 32, "tagstack::slice::Slice layout");

__device__ __forceinline__ uint64_t mix64(uint64_t x) {
  x ^= x >> 33;
  x *= 0xff51afd7ed558ccdull;
  x ^= x >> 33;
  x *= 0xc4ceb9fe1a85ec53ull;
  x ^= x >> 33;
  return x;
}

struct JoinArgs {
  const hbt_slice* slices;
  uint64_t n_slices;
  uint64_t offset;             // consumer mark: slices before it were consumed earlier
  const uint64_t* mark_before; // pass 2: [n_intervals] mark in front of interval k (from pass 1), or null
  uint64_t* mark_after;        // pass 1: [n_intervals] mark after interval k if it hands out slices, else 0
  const uint64_t* intervals;   // interval k: {tstamp, duration} at intervals + k * stride
  uint32_t stride_words;
  uint64_t n_intervals;
  int trim, abort_on_missing;
  uint64_t* first;             // [n_intervals] first included slice
  uint64_t* counts;            // [n_intervals + 1] included slices (0 = the reference returns nullptr)
  unsigned long long* mark;    // consumer mark after the last interval that consumed something
  uint32_t* error;             // 1: an interval of zero duration (HBT_ARG_CHECK_GT)
};

// first index in [lo, hi) whose slice does NOT satisfy pred (pred is true on a prefix)
template <class P>
__device__ __forceinline__ uint64_t partition_point(uint64_t lo, uint64_t hi, P pred) {
  while (lo < hi) {
    const uint64_t mid = lo + ((hi - lo) >> 1);
    if (pred(mid)) lo = mid + 1;
    else hi = mid;
  }
  return lo;
}

// One warp per interval: which slices makeIntervalIntersectingSlices returns for it.
//
// The consumer mark.  The reference starts every interval's walk at a mark that the earlier
// intervals advanced by what they consumed.  A slice consumed by an earlier interval ends at or
// before this interval's start, so the mark matters in exactly one case: a slice that ends EXACTLY
// where the interval starts is still walked into (its end is not < the start, :79) unless an earlier
// interval consumed it.  Which intervals hand out slices does not depend on that case in a way that
// changes the marks (where it flips the outcome, the mark is already past the slice), so the marks
// are a running maximum: pass 1 computes every interval's mark-after from the initial mark alone,
// an exclusive max-scan turns them into marks-before, pass 2 decides with those.
__global__ void __launch_bounds__(256) k_join_plan(const JoinArgs a) {
  const uint32_t lane = threadIdx.x & 31;
  const uint64_t warp0 = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) >> 5;
  const uint64_t n_warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  for (uint64_t k = warp0; k < a.n_intervals; k += n_warps) {
    const uint64_t tstamp = a.intervals[k * a.stride_words], duration = a.intervals[k * a.stride_words + 1];
    const uint64_t t_end = tstamp + duration;
    if (duration == 0 && lane == 0) atomicOr(a.error, 1u);
    const hbt_slice* s = a.slices;
    // the walk starts at the consumer mark, which earlier intervals may have moved (see below)
    uint64_t from = a.offset;
    if (a.mark_before && a.mark_before[k] > from) from = a.mark_before[k];
    // "move until the first slice that ends after count has started" (:79-84): skip while end < tstamp
    const uint64_t f = partition_point(from, a.n_slices, [&](uint64_t i) { return s[i].tstamp + s[i].duration < tstamp; });
    // "add slices until last one or end of time interval" (:104): take while tstamp < t_end
    uint64_t g = partition_point(f, a.n_slices, [&](uint64_t i) { return s[i].tstamp < t_end; });
    bool ok = f < a.n_slices && duration != 0;           // no slice left: {nullptr, 0} (:87-91)
    if (ok && a.abort_on_missing && s[f].tstamp > tstamp) ok = false;  // :93-97
    if (ok && !a.trim && g > f && s[g - 1].tstamp + s[g - 1].duration > t_end) --g;  // :113-115 (only the last can cross)
    if (ok && a.abort_on_missing) {
      // a hole between two consecutive slices aborts (:107-111): checked by all lanes
      bool hole = false;
      for (uint64_t i = f + 1 + lane; i < g; i += 32) hole |= s[i - 1].tstamp + s[i - 1].duration != s[i].tstamp;
      if (__any_sync(0xffffffffu, hole)) ok = false;
    }
    if (ok && g == f) ok = false;                        // nothing intersects (:129-134)
    if (ok && a.abort_on_missing && s[g - 1].tstamp + s[g - 1].duration < t_end) ok = false;  // :139-145
    if (ok && !a.mark_after) {
      // what the reference asserts on before returning (:179-181): the slices handed out may not be
      // longer than the interval -- without trimming, a slice reaching in from before the start makes
      // them so, and makeIntervalIntersectingSlices throws
      unsigned long long sum = 0;
      for (uint64_t i = f + lane; i < g; i += 32) {
        uint64_t b = s[i].tstamp, e = b + s[i].duration;
        if (a.trim) {
          if (b < tstamp) b = tstamp;
          if (e > t_end) e = t_end;
        }
        sum += e - b;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
      if (sum > duration && lane == 0) atomicOr(a.error, 32u);
    }
    if (lane == 0) {
      // consumed = skipped + included, minus the last one when it crosses the interval's end (:164-167)
      const uint64_t after = ok ? g - ((s[g - 1].tstamp + s[g - 1].duration > t_end) ? 1 : 0) : 0;
      if (a.mark_after) {
        a.mark_after[k] = after;
      } else {
        a.first[k] = f;
        a.counts[k] = ok ? g - f : 0;
        if (ok) atomicMax(a.mark, (unsigned long long)after);  // (marks grow with the interval index)
      }
    }
  }
}

// One warp per interval: copy its slices to the output, trimming the two boundary slices (:147-177).
__global__ void __launch_bounds__(256)
    k_join_emit(const JoinArgs a, const uint64_t* __restrict__ out_offsets, hbt_slice* __restrict__ out, uint64_t cap) {
  const uint32_t lane = threadIdx.x & 31;
  const uint64_t warp0 = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) >> 5;
  const uint64_t n_warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  for (uint64_t k = warp0; k < a.n_intervals; k += n_warps) {
    const uint64_t o0 = out_offsets[k], n = out_offsets[k + 1] - o0;
    if (n == 0 || o0 + n > cap) continue;
    const uint64_t tstamp = a.intervals[k * a.stride_words], t_end = tstamp + a.intervals[k * a.stride_words + 1];
    const uint64_t f = a.first[k];
    for (uint64_t i = lane; i < n; i += 32) {
      hbt_slice s = a.slices[f + i];
      if (a.trim) {
        const uint64_t end = s.tstamp + s.duration;
        if (i == 0 && s.tstamp < tstamp) {  // trim the start of the first slice
          s.duration = end - tstamp;
          s.tstamp = tstamp;
          s.swin_type = kAnalysis;
        }
        if (i == n - 1 && end > t_end) {    // the last slice goes beyond the interval
          s.duration -= end - t_end;
          s.swout_type = kAnalysis;
        }
      }
      out[o0 + i] = s;
    }
  }
}

// exclusive prefix scan of v[0..n) in place (v has n + 1 entries; v[n] receives the total), with
// + or max as the operator (identity 0 for both: the values are unsigned): per-block scans, one
// block scans the block totals, the blocks fold their base in.
constexpr int kScanBlock = 1024;
template <bool MAX>
__device__ __forceinline__ uint64_t scan_op(uint64_t a, uint64_t b) {
  return MAX ? (a > b ? a : b) : a + b;
}
template <bool MAX>
__device__ __forceinline__ void block_inclusive_scan(uint64_t* s) {
  for (int o = 1; o < kScanBlock; o <<= 1) {
    const uint64_t y = threadIdx.x >= (unsigned)o ? s[threadIdx.x - o] : 0;
    __syncthreads();
    s[threadIdx.x] = scan_op<MAX>(s[threadIdx.x], y);
    __syncthreads();
  }
}
template <bool MAX>
__global__ void __launch_bounds__(kScanBlock) k_scan_blocks(uint64_t* v, uint64_t n, uint64_t* block_sums) {
  __shared__ uint64_t s[kScanBlock];
  const uint64_t i = blockIdx.x * (uint64_t)kScanBlock + threadIdx.x;
  s[threadIdx.x] = i < n ? v[i] : 0;
  __syncthreads();
  block_inclusive_scan<MAX>(s);
  if (i < n) v[i] = threadIdx.x ? s[threadIdx.x - 1] : 0;
  if (threadIdx.x == kScanBlock - 1) block_sums[blockIdx.x] = s[threadIdx.x];
}
template <bool MAX>
__global__ void __launch_bounds__(kScanBlock) k_scan_sums(uint64_t* block_sums, uint32_t n_blocks, uint64_t* total) {
  __shared__ uint64_t carry;
  __shared__ uint64_t s[kScanBlock];
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (uint32_t b0 = 0; b0 < n_blocks; b0 += kScanBlock) {
    const uint32_t i = b0 + threadIdx.x;
    s[threadIdx.x] = i < n_blocks ? block_sums[i] : 0;
    __syncthreads();
    block_inclusive_scan<MAX>(s);
    const uint64_t excl = scan_op<MAX>(carry, threadIdx.x ? s[threadIdx.x - 1] : 0);
    const uint64_t last = scan_op<MAX>(carry, s[kScanBlock - 1]);
    __syncthreads();
    if (i < n_blocks) block_sums[i] = excl;
    if (threadIdx.x == 0) carry = last;
    __syncthreads();
  }
  if (threadIdx.x == 0) *total = carry;
}
template <bool MAX>
__global__ void __launch_bounds__(kScanBlock) k_scan_add(uint64_t* v, uint64_t n, const uint64_t* block_sums, const uint64_t* total) {
  const uint64_t i = blockIdx.x * (uint64_t)kScanBlock + threadIdx.x;
  if (i < n) v[i] = scan_op<MAX>(v[i], block_sums[blockIdx.x]);
  if (i == 0) v[n] = *total;
}

// ---- frequencies -----------------------------------------------------------------------------
struct FreqTable {
  uint64_t* keys;       // [cap + 1]; slot cap is kInvalidTagStackId's
  unsigned long long* duration;
  unsigned long long* num_obs;
  unsigned long long* num_intervals;
  uint32_t* bid;        // bin of the stack id (TagStackIdBinner)
  uint64_t cap;         // power of two
  uint32_t* error;      // 2: table full, 4: pair table full, 8: a stack id the bins do not know
};

__device__ __forceinline__ uint64_t freq_slot(const FreqTable& t, uint64_t sid, bool insert) {
  if (sid == kEmptyKey) return t.cap;
  uint64_t h = mix64(sid) & (t.cap - 1);
  for (uint64_t probes = 0; probes < t.cap; ++probes) {
    uint64_t k = t.keys[h];
    if (k == sid) return h;
    if (k == kEmptyKey) {
      if (!insert) return kEmptyKey;
      k = atomicCAS(reinterpret_cast<unsigned long long*>(t.keys + h), (unsigned long long)kEmptyKey,
                    (unsigned long long)sid);
      if (k == kEmptyKey || k == sid) return h;
    }
    h = (h + 1) & (t.cap - 1);
  }
  atomicOr(t.error, 2u);
  return kEmptyKey;
}

// One warp per interval (CuMonData::stackIdFreqs, MonData.h:245-266): every slice adds its duration
// and one observation to its stack id; the first slice of a stack id WITHIN the interval adds one
// interval -- decided by who inserts the pair (slot, interval) into the pair table.
__global__ void __launch_bounds__(256)
    k_freqs_add(FreqTable t, const hbt_slice* __restrict__ sl, const uint64_t* __restrict__ offsets, uint64_t n_intervals,
                unsigned long long* pairs, uint64_t pair_cap) {
  const uint32_t lane = threadIdx.x & 31;
  const uint64_t warp0 = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) >> 5;
  const uint64_t n_warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  for (uint64_t k = warp0; k < n_intervals; k += n_warps) {
    for (uint64_t i = offsets[k] + lane; i < offsets[k + 1]; i += 32) {
      const hbt_slice s = sl[i];
      const uint64_t slot = freq_slot(t, s.stack_id, true);
      if (slot == kEmptyKey) continue;
      atomicAdd(t.num_obs + slot, 1ull);
      atomicAdd(t.duration + slot, (unsigned long long)s.duration);
      // pair key: slot and interval, both < 2^32 (checked by the host)
      const unsigned long long pk = ((unsigned long long)slot << 32) | (unsigned long long)k;
      uint64_t h = mix64(pk) & (pair_cap - 1);
      for (uint64_t probes = 0;; ++probes) {
        if (probes == pair_cap) {
          atomicOr(t.error, 4u);
          break;
        }
        const unsigned long long old = atomicCAS(pairs + h, ~0ull, pk);
        if (old == ~0ull) {  // first occurrence of this stack id in this interval
          atomicAdd(t.num_intervals + slot, 1ull);
          break;
        }
        if (old == pk) break;
        h = (h + 1) & (pair_cap - 1);
      }
    }
  }
}

// TagStackIdBinner's matrix fill (MonData.h:563-606) for one compute unit.  row_of[k]: row of
// interval k = its rank among this unit's intervals with a non-zero total duration, or ~0.
__global__ void __launch_bounds__(256)
    k_bin_rows(const hbt_slice* __restrict__ sl, const uint64_t* __restrict__ offsets, uint64_t n_intervals,
               uint64_t* row_flag) {
  const uint32_t lane = threadIdx.x & 31;
  const uint64_t warp0 = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) >> 5;
  const uint64_t n_warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  for (uint64_t k = warp0; k < n_intervals; k += n_warps) {
    bool any = false;
    for (uint64_t i = offsets[k] + lane; i < offsets[k + 1]; i += 32) any |= sl[i].duration != 0;
    any = __any_sync(0xffffffffu, any);
    if (lane == 0) row_flag[k] = any ? 1 : 0;  // "skip rows with zero durations" (:597-602)
  }
}
__global__ void __launch_bounds__(256)
    k_bin_add(FreqTable t, const hbt_slice* __restrict__ sl, const uint64_t* __restrict__ offsets, uint64_t n_intervals,
              const uint64_t* __restrict__ row_of, unsigned long long* matrix, uint64_t n_rows, uint32_t n_bins) {
  const uint32_t lane = threadIdx.x & 31;
  const uint64_t warp0 = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) >> 5;
  const uint64_t n_warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  for (uint64_t k = warp0; k < n_intervals; k += n_warps) {
    if (row_of[k + 1] == row_of[k]) continue;   // an interval without duration takes no row
    const uint64_t row = row_of[k];
    if (row >= n_rows) {
      if (lane == 0) atomicOr(t.error, 16u);
      continue;
    }
    for (uint64_t i = offsets[k] + lane; i < offsets[k + 1]; i += 32) {
      const hbt_slice s = sl[i];
      if (s.duration == 0) continue;            // zero-duration slices are skipped with a warning (:584-592)
      const uint64_t slot = freq_slot(t, s.stack_id, false);
      if (slot == kEmptyKey) {
        atomicOr(t.error, 8u);                  // binkey_to_bid.at() would throw
        continue;
      }
      atomicAdd(matrix + row * n_bins + t.bid[slot], (unsigned long long)s.duration);
    }
  }
}

}  // namespace

struct hbt_slice_stats {
  hbt_ctx* ctx = nullptr;
  uint64_t cap = 0;                 // stack-id slots (power of two) + 1
  uint64_t* d_keys = nullptr;
  unsigned long long *d_duration = nullptr, *d_obs = nullptr, *d_nint = nullptr;
  uint32_t* d_bid = nullptr;
  uint32_t* d_error = nullptr;
  unsigned long long* d_pairs = nullptr;
  uint64_t pair_cap = 0;
  // scratch of the join / the matrix
  uint64_t *d_first = nullptr, *d_scan = nullptr, *d_block_sums = nullptr;
  unsigned long long* d_mark = nullptr;
  uint64_t cap_intervals = 0;
  unsigned long long* d_matrix = nullptr;
  uint64_t n_rows = 0;
  uint32_t n_bins = 0;
};

namespace {

#define SCU(call)                                                                                     \
  do {                                                                                                \
    cudaError_t e__ = (call);                                                                         \
    if (e__ != cudaSuccess)                                                                           \
      return hbt_ctx_fail(st->ctx, e__ == cudaErrorMemoryAllocation ? -ENOMEM : -EIO, "%s: %s", #call, \
                          cudaGetErrorString(e__));                                                   \
  } while (0)

FreqTable table_of(const hbt_slice_stats* st) {
  return FreqTable{st->d_keys, st->d_duration, st->d_obs, st->d_nint, st->d_bid, st->cap, st->d_error};
}

int grid_for(const hbt_slice_stats* st, uint64_t warps) {
  uint64_t g = (warps + 7) / 8;
  const uint64_t cap = (uint64_t)hbt_ctx_sm_count(st->ctx) * 8;
  if (g > cap) g = cap;
  return (int)(g ? g : 1);
}

int reserve_intervals(hbt_slice_stats* st, uint64_t n) {
  if (n <= st->cap_intervals) return 0;
  if (st->d_first) cudaFree(st->d_first);
  if (st->d_scan) cudaFree(st->d_scan);
  if (st->d_block_sums) cudaFree(st->d_block_sums);
  st->d_first = st->d_scan = st->d_block_sums = nullptr;
  st->cap_intervals = 0;
  SCU(cudaMalloc(&st->d_first, n * 8));
  SCU(cudaMalloc(&st->d_scan, (n + 1) * 8));
  SCU(cudaMalloc(&st->d_block_sums, ((n + kScanBlock - 1) / kScanBlock + 1) * 8));
  st->cap_intervals = n;
  return 0;
}

// v[0..n) -> exclusive prefix sums, v[n] = total (device); returns the total on the host
template <bool MAX>
int scan_inplace(hbt_slice_stats* st, uint64_t* v, uint64_t n, uint64_t* total_host) {
  cudaStream_t s = hbt_ctx_stream(st->ctx);
  const uint32_t nb = (uint32_t)((n + kScanBlock - 1) / kScanBlock);
  uint64_t* d_total = st->d_block_sums + nb;
  k_scan_blocks<MAX><<<nb, kScanBlock, 0, s>>>(v, n, st->d_block_sums);
  k_scan_sums<MAX><<<1, kScanBlock, 0, s>>>(st->d_block_sums, nb, d_total);
  k_scan_add<MAX><<<nb, kScanBlock, 0, s>>>(v, n, st->d_block_sums, d_total);
  hbt_ctx_count_launches(st->ctx, 3);
  SCU(cudaGetLastError());
  if (total_host) {
    SCU(cudaMemcpyAsync(total_host, d_total, 8, cudaMemcpyDeviceToHost, s));
    SCU(cudaStreamSynchronize(s));
  }
  return 0;
}

int check_error(hbt_slice_stats* st) {
  uint32_t e = 0;
  cudaStream_t s = hbt_ctx_stream(st->ctx);
  SCU(cudaMemcpyAsync(&e, st->d_error, 4, cudaMemcpyDeviceToHost, s));
  SCU(cudaStreamSynchronize(s));
  if (!e) return 0;
  SCU(cudaMemsetAsync(st->d_error, 0, 4, s));
  if (e & 1) return hbt_ctx_fail(st->ctx, -EINVAL, "an interval of zero duration");
  if (e & 32)
    return hbt_ctx_fail(st->ctx, -EDOM, "the slices of an interval are longer than the interval (untrimmed slices "
                                        "reaching across its start): the reference throws here");
  if (e & 2) return hbt_ctx_fail(st->ctx, -ENOSPC, "more distinct stack ids than hbt_slice_stats_create was given");
  if (e & 8) return hbt_ctx_fail(st->ctx, -ENOENT, "a slice whose stack id was not in the frequencies the bins were made from");
  if (e & 16) return hbt_ctx_fail(st->ctx, -ERANGE, "more rows than the matrix was made for");
  return hbt_ctx_fail(st->ctx, -EIO, "pair table full");
}

}  // namespace

extern "C" {

int hbt_slice_stats_create(hbt_ctx* ctx, hbt_slice_stats** out, uint64_t max_stack_ids) {
  if (!ctx || !out) return -EINVAL;
  *out = nullptr;
  if (max_stack_ids == 0 || max_stack_ids > (1ull << 30)) return hbt_ctx_fail(ctx, -EINVAL, "max_stack_ids out of range");
  hbt_slice_stats* st = new hbt_slice_stats();
  st->ctx = ctx;
  uint64_t cap = 64;
  while (cap < 2 * max_stack_ids) cap <<= 1;  // load <= 1/2
  st->cap = cap;
  cudaSetDevice(hbt_ctx_device(ctx));
  const size_t n = cap + 1;
  bool ok = cudaMalloc(&st->d_keys, n * 8) == cudaSuccess && cudaMalloc(&st->d_duration, n * 8) == cudaSuccess &&
            cudaMalloc(&st->d_obs, n * 8) == cudaSuccess && cudaMalloc(&st->d_nint, n * 8) == cudaSuccess &&
            cudaMalloc(&st->d_bid, n * 4) == cudaSuccess && cudaMalloc(&st->d_error, 4) == cudaSuccess &&
            cudaMalloc(&st->d_mark, 8) == cudaSuccess;
  if (!ok) {
    cudaGetLastError();
    hbt_slice_stats_destroy(st);
    return hbt_ctx_fail(ctx, -ENOMEM, "hbt_slice_stats_create: out of device memory");
  }
  *out = st;
  return hbt_slice_stats_clear(st);
}

void hbt_slice_stats_destroy(hbt_slice_stats* st) {
  if (!st) return;
  cudaSetDevice(hbt_ctx_device(st->ctx));
  cudaStreamSynchronize(hbt_ctx_stream(st->ctx));
  void* ptrs[] = {st->d_keys, st->d_duration, st->d_obs, st->d_nint, st->d_bid, st->d_error, st->d_pairs,
                  st->d_first, st->d_scan, st->d_block_sums, st->d_mark, st->d_matrix};
  for (void* p : ptrs)
    if (p) cudaFree(p);
  delete st;
}

int hbt_slice_stats_clear(hbt_slice_stats* st) {
  if (!st) return -EINVAL;
  cudaStream_t s = hbt_ctx_stream(st->ctx);
  SCU(cudaSetDevice(hbt_ctx_device(st->ctx)));
  const size_t n = st->cap + 1;
  SCU(cudaMemsetAsync(st->d_keys, 0xFF, n * 8, s));
  SCU(cudaMemsetAsync(st->d_duration, 0, n * 8, s));
  SCU(cudaMemsetAsync(st->d_obs, 0, n * 8, s));
  SCU(cudaMemsetAsync(st->d_nint, 0, n * 8, s));
  SCU(cudaMemsetAsync(st->d_bid, 0, n * 4, s));
  SCU(cudaMemsetAsync(st->d_error, 0, 4, s));
  return 0;
}

int64_t hbt_slices_join_dev(hbt_slice_stats* st, const hbt_slice* slices, uint64_t n_slices, uint64_t offset,
                            const uint64_t* intervals, uint32_t interval_stride_words, uint64_t n_intervals,
                            int trim, int abort_on_missing_slices, hbt_slice* out, uint64_t cap,
                            uint64_t* out_offsets, uint64_t* new_offset) {
  if (!st) return -EINVAL;
  if (!out_offsets || (n_intervals && (!intervals || interval_stride_words < 2)) || (n_slices && !slices))
    return hbt_ctx_fail(st->ctx, -EINVAL, "bad arguments");
  if (offset > n_slices) return hbt_ctx_fail(st->ctx, -EINVAL, "offset beyond the slices");
  cudaStream_t s = hbt_ctx_stream(st->ctx);
  SCU(cudaSetDevice(hbt_ctx_device(st->ctx)));
  if (new_offset) *new_offset = offset;
  if (n_intervals == 0) {
    SCU(cudaMemsetAsync(out_offsets, 0, 8, s));
    return 0;
  }
  if (int rc = reserve_intervals(st, n_intervals)) return rc;
  const unsigned long long mark0 = offset;
  SCU(cudaMemcpyAsync(st->d_mark, &mark0, 8, cudaMemcpyHostToDevice, s));
  JoinArgs a{};
  a.slices = slices;
  a.n_slices = n_slices;
  a.offset = offset;
  a.intervals = intervals;
  a.stride_words = interval_stride_words;
  a.n_intervals = n_intervals;
  a.trim = trim;
  a.abort_on_missing = abort_on_missing_slices;
  a.first = st->d_first;
  a.counts = out_offsets;   // counted in place, then scanned into offsets
  a.mark = st->d_mark;
  a.error = st->d_error;
  // pass 1: marks after every interval from the initial mark; exclusive max-scan -> marks before
  a.mark_after = st->d_scan;
  k_join_plan<<<grid_for(st, n_intervals), 256, 0, s>>>(a);
  hbt_ctx_count_launches(st->ctx, 1);
  SCU(cudaGetLastError());
  if (int rc = scan_inplace<true>(st, st->d_scan, n_intervals, nullptr)) return rc;
  // pass 2: the plan proper
  a.mark_after = nullptr;
  a.mark_before = st->d_scan;
  k_join_plan<<<grid_for(st, n_intervals), 256, 0, s>>>(a);
  hbt_ctx_count_launches(st->ctx, 1);
  SCU(cudaGetLastError());
  uint64_t total = 0;
  if (int rc = scan_inplace<false>(st, out_offsets, n_intervals, &total)) return rc;
  if (int rc = check_error(st)) return rc;
  if (total > cap || (total && !out))
    return hbt_ctx_fail(st->ctx, -ENOSPC, "cap %llu < %llu slices", (unsigned long long)cap, (unsigned long long)total);
  if (total) {
    k_join_emit<<<grid_for(st, n_intervals), 256, 0, s>>>(a, out_offsets, out, cap);
    hbt_ctx_count_launches(st->ctx, 1);
    SCU(cudaGetLastError());
  }
  if (new_offset) {
    unsigned long long m = 0;
    SCU(cudaMemcpyAsync(&m, st->d_mark, 8, cudaMemcpyDeviceToHost, s));
    SCU(cudaStreamSynchronize(s));
    *new_offset = m;
  }
  return (int64_t)total;
}

int hbt_slice_stats_add_unit_dev(hbt_slice_stats* st, const hbt_slice* pi_slices, const uint64_t* offsets,
                                 uint64_t n_intervals, uint64_t n_slices) {
  if (!st || (n_intervals && !offsets) || (n_slices && !pi_slices)) return -EINVAL;
  if (n_intervals == 0 || n_slices == 0) return 0;
  if (n_intervals >= (1ull << 32)) return hbt_ctx_fail(st->ctx, -EINVAL, "too many intervals in one call");
  cudaStream_t s = hbt_ctx_stream(st->ctx);
  SCU(cudaSetDevice(hbt_ctx_device(st->ctx)));
  uint64_t pc = 64;
  while (pc < 2 * n_slices) pc <<= 1;
  if (pc > st->pair_cap) {
    if (st->d_pairs) cudaFree(st->d_pairs);
    st->d_pairs = nullptr;
    st->pair_cap = 0;
    SCU(cudaMalloc(&st->d_pairs, pc * 8));
    st->pair_cap = pc;
  }
  // (the pairs are per compute unit: interval indices of different units are different intervals)
  SCU(cudaMemsetAsync(st->d_pairs, 0xFF, st->pair_cap * 8, s));
  k_freqs_add<<<grid_for(st, n_intervals), 256, 0, s>>>(table_of(st), pi_slices, offsets, n_intervals, st->d_pairs,
                                                        st->pair_cap);
  hbt_ctx_count_launches(st->ctx, 1);
  SCU(cudaGetLastError());
  return check_error(st);
}

int64_t hbt_slice_stats_fetch(hbt_slice_stats* st, hbt_slice_freq* out, uint64_t cap) {
  if (!st) return -EINVAL;
  cudaStream_t s = hbt_ctx_stream(st->ctx);
  SCU(cudaSetDevice(hbt_ctx_device(st->ctx)));
  const size_t n = st->cap + 1;
  std::vector<uint64_t> keys(n), dur(n), obs(n), nint(n);
  SCU(cudaMemcpyAsync(keys.data(), st->d_keys, n * 8, cudaMemcpyDeviceToHost, s));
  SCU(cudaMemcpyAsync(dur.data(), st->d_duration, n * 8, cudaMemcpyDeviceToHost, s));
  SCU(cudaMemcpyAsync(obs.data(), st->d_obs, n * 8, cudaMemcpyDeviceToHost, s));
  SCU(cudaMemcpyAsync(nint.data(), st->d_nint, n * 8, cudaMemcpyDeviceToHost, s));
  SCU(cudaStreamSynchronize(s));
  std::vector<hbt_slice_freq> v;
  for (size_t i = 0; i < n; ++i) {
    const bool used = i == st->cap ? obs[i] != 0 : keys[i] != kEmptyKey;
    if (used) v.push_back(hbt_slice_freq{i == st->cap ? kEmptyKey : keys[i], dur[i], obs[i], nint[i]});
  }
  std::sort(v.begin(), v.end(), [](const hbt_slice_freq& a, const hbt_slice_freq& b) { return a.stack_id < b.stack_id; });
  if (out) {
    if (v.size() > cap) return hbt_ctx_fail(st->ctx, -ENOSPC, "cap %llu < %zu stack ids", (unsigned long long)cap, v.size());
    if (!v.empty()) memcpy(out, v.data(), v.size() * sizeof(hbt_slice_freq));
  }
  return (int64_t)v.size();
}

int64_t hbt_slice_bins_prepare(hbt_slice_stats* st, uint64_t min_distinct_intervals, uint64_t min_duration,
                               uint64_t n_rows, uint64_t* stack_ids, uint32_t* bids, uint64_t cap) {
  if (!st) return -EINVAL;
  cudaStream_t s = hbt_ctx_stream(st->ctx);
  SCU(cudaSetDevice(hbt_ctx_device(st->ctx)));
  const size_t n = st->cap + 1;
  std::vector<uint64_t> keys(n), dur(n), obs(n), nint(n);
  SCU(cudaMemcpyAsync(keys.data(), st->d_keys, n * 8, cudaMemcpyDeviceToHost, s));
  SCU(cudaMemcpyAsync(dur.data(), st->d_duration, n * 8, cudaMemcpyDeviceToHost, s));
  SCU(cudaMemcpyAsync(obs.data(), st->d_obs, n * 8, cudaMemcpyDeviceToHost, s));
  SCU(cudaMemcpyAsync(nint.data(), st->d_nint, n * 8, cudaMemcpyDeviceToHost, s));
  SCU(cudaStreamSynchronize(s));
  // bin 0 collects the ignored stack ids (ignored_bid = reserveBinId(), :533); the others get their own
  // bin, numbered here in ascending stack-id order (the reference numbers them in the iteration order of
  // an unordered_map, which is not defined: compare matrices through stack ids, not raw bin ids)
  std::vector<std::pair<uint64_t, size_t>> used;
  for (size_t i = 0; i < n; ++i)
    if (i == st->cap ? obs[i] != 0 : keys[i] != kEmptyKey) used.emplace_back(i == st->cap ? kEmptyKey : keys[i], i);
  std::sort(used.begin(), used.end());
  std::vector<uint32_t> bid(n, 0);
  uint32_t next = 1;
  if (stack_ids && used.size() > cap) return hbt_ctx_fail(st->ctx, -ENOSPC, "cap %llu < %zu stack ids", (unsigned long long)cap, used.size());
  for (size_t j = 0; j < used.size(); ++j) {
    const size_t i = used[j].second;
    const uint32_t b = (nint[i] >= min_distinct_intervals && dur[i] >= min_duration) ? next++ : 0;
    bid[i] = b;
    if (stack_ids) stack_ids[j] = used[j].first;
    if (bids) bids[j] = b;
  }
  SCU(cudaMemcpyAsync(st->d_bid, bid.data(), n * 4, cudaMemcpyHostToDevice, s));
  if (st->d_matrix) cudaFree(st->d_matrix);
  st->d_matrix = nullptr;
  st->n_rows = n_rows;
  st->n_bins = next;
  const size_t cells = (size_t)n_rows * next;
  SCU(cudaMalloc(&st->d_matrix, (cells ? cells : 1) * 8));
  SCU(cudaMemsetAsync(st->d_matrix, 0, (cells ? cells : 1) * 8, s));
  SCU(cudaStreamSynchronize(s));  // bid[] is a local
  return (int64_t)next;
}

int hbt_slice_bins_add_unit_dev(hbt_slice_stats* st, const hbt_slice* pi_slices, const uint64_t* offsets,
                                uint64_t n_intervals) {
  if (!st || !st->d_matrix) return st ? hbt_ctx_fail(st->ctx, -EINVAL, "hbt_slice_bins_prepare first") : -EINVAL;
  if (n_intervals == 0) return 0;
  if (!offsets) return -EINVAL;
  cudaStream_t s = hbt_ctx_stream(st->ctx);
  SCU(cudaSetDevice(hbt_ctx_device(st->ctx)));
  if (int rc = reserve_intervals(st, n_intervals)) return rc;
  k_bin_rows<<<grid_for(st, n_intervals), 256, 0, s>>>(pi_slices, offsets, n_intervals, st->d_scan);
  hbt_ctx_count_launches(st->ctx, 1);
  uint64_t rows_used = 0;
  if (int rc = scan_inplace<false>(st, st->d_scan, n_intervals, &rows_used)) return rc;
  k_bin_add<<<grid_for(st, n_intervals), 256, 0, s>>>(table_of(st), pi_slices, offsets, n_intervals, st->d_scan,
                                                      st->d_matrix, st->n_rows, st->n_bins);
  hbt_ctx_count_launches(st->ctx, 1);
  SCU(cudaGetLastError());
  return check_error(st);
}

int hbt_slice_bins_fetch(hbt_slice_stats* st, uint64_t* matrix, uint64_t cap_cells) {
  if (!st || !matrix || !st->d_matrix) return -EINVAL;
  const uint64_t cells = st->n_rows * st->n_bins;
  if (cells > cap_cells) return hbt_ctx_fail(st->ctx, -ENOSPC, "cap %llu < %llu cells", (unsigned long long)cap_cells, (unsigned long long)cells);
  cudaStream_t s = hbt_ctx_stream(st->ctx);
  SCU(cudaSetDevice(hbt_ctx_device(st->ctx)));
  if (cells) SCU(cudaMemcpyAsync(matrix, st->d_matrix, cells * 8, cudaMemcpyDeviceToHost, s));
  SCU(cudaStreamSynchronize(s));
  return 0;
}

}  // extern "C"
