// Counter-group kernels (K4 group reduce, K5 scale + derive).
//
// Replaces, for a recorded trace of GroupReadValues<TMode>::T
// (hbt/src/perf_event/CpuEventsGroup.h:371-379):
//   PerCpuBase<T>::read           hbt/src/perf_event/PerCpuBase.h:104-119
//   GroupReadValues::zero/accum   CpuEventsGroup.h:508-522
//   GroupReadValues::getCount     CpuEventsGroup.h:449-459
//   PerfMonitor::log rate math    dynolog/src/PerfMonitor.cpp:52-67
//
// Layout in HBM: reads[interval][cpu][3+E] u64, exactly the kernel's read_format
// (PERF_FORMAT_GROUP | TOTAL_TIME_ENABLED | TOTAL_TIME_RUNNING) per (interval,
// cpu).  The reduce is a pure streaming sum: every input word is read once
// (8*(3+E) B per (interval,cpu)) and 8*(3+E) B are written per interval, so the
// kernel is HBM-bound.  Integer sums wrap mod 2^64 like the reference's u64 adds;
// the scaling is done AFTER the sum over CPUs (sum-of-scaled != scaled-of-sum).
#pragma once
#include <cstdint>
#include "../../include/hbt_b200.h"
#include "ptx.cuh"

namespace hbt {

// GroupReadValues::getCount (CpuEventsGroup.h:449-459): IEEE double mul then
// div, round-to-nearest, truncating conversion; 0 when time_enabled == 0.
// time_running == 0 with time_enabled != 0 is undefined in the reference
// (division by zero cast to u64); here it saturates (documented in DESIGN.md).
__device__ __forceinline__ uint64_t scaled_count(uint64_t count, uint64_t te, uint64_t tr) {
  if (te == 0) return 0;
  const double v =
      __ddiv_rn(__dmul_rn(__ull2double_rn(count), __ull2double_rn(te)), __ull2double_rn(tr));
  return __double2ull_rz(v);
}

// PerfMonitor.cpp:56-57: float(double(count) * 1e3 / double(time_running)).
__device__ __forceinline__ float rate_per_ms(uint64_t count, uint64_t tr) {
  return __double2float_rn(
      __ddiv_rn(__dmul_rn(__ull2double_rn(count), 1e3), __ull2double_rn(tr)));
}

__device__ __forceinline__ double ratio_of(uint64_t a, uint64_t b) {
  return __ddiv_rn(__ull2double_rn(a), __ull2double_rn(b));
}

struct CountsArgs {
  const uint64_t* reads;  // [I][C][W]
  uint64_t n_intervals;
  uint32_t n_cpus;
  uint32_t n_events;
  const hbt_ratio* ratios;  // device copy
  uint32_t n_ratios;
  uint64_t* sums;      // [I][W] or null
  uint64_t* scaled;    // [I][E] or null
  float* rates;        // [I][E] or null
  double* ratio_out;   // [I][R] or null
  int do_scale;        // 0: K4 only
};

// One WARP per interval at a time (grid-stride over intervals), no shared memory and
// no block barrier: the 64 resident warps of an SM each stream their own interval.
//  - W = 3+E <= 32: the first `active = floor(32/W)*W` lanes read words lane,
//    lane+active, ... of the interval's contiguous C*W words, so a lane's column
//    (lane % W) is fixed and its partial sum lives in one register; a warp request is
//    `active` consecutive u64s.  Lanes with equal column are then combined by shuffles.
//  - W > 32 (WIDE): lane l owns columns l, l+32, l+64 of every CPU's row.
// The scale/derive epilogue runs in the same warp (FP64 on lanes that hold a count).
template <bool WIDE>
__global__ void __launch_bounds__(256) k_counts_reduce(const CountsArgs a) {
  const uint32_t E = a.n_events;
  const uint32_t W = 3 + E;
  const uint32_t lane = threadIdx.x & 31;
  const uint64_t warp0 = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) >> 5;
  const uint64_t n_warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  const uint64_t N = (uint64_t)a.n_cpus * W;
  constexpr unsigned FULL = 0xffffffffu;

  for (uint64_t i = warp0; i < a.n_intervals; i += n_warps) {
    const uint64_t* base = a.reads + i * N;
    // s0/s1/s2: the sums of columns lane, lane+32, lane+64 (column 0 counts failed reads)
    uint64_t s0 = 0, s1 = 0, s2 = 0;
    if (!WIDE) {
      const uint32_t groups = 32 / W, active = groups * W;
      const uint32_t col = lane % W;
      uint64_t acc = 0;
      if (lane < active) {
        // word0 = n_events of each T; a mismatch marks a failed read
        // (CpuEventsGroupBase::read, CpuEventsGroup.h:599-614).  One loop for all
        // columns (a select, not a branch: the lanes of a warp must not diverge here).
        const bool is0 = col == 0;
#pragma unroll 8
        for (uint64_t f = lane; f < N; f += active) {
          const uint64_t v = ld_stream_u64(base + f);
          acc += is0 ? (uint64_t)(v != E) : v;
        }
      }
      s0 = acc;
      for (uint32_t k = 1; k < groups; ++k) {
        const uint64_t o = __shfl_sync(FULL, acc, (lane + k * W) & 31);
        if (lane < W) s0 += o;
      }
    } else {
      const bool has1 = lane + 32 < W, has2 = lane + 64 < W;
#pragma unroll 4
      for (uint32_t c = 0; c < a.n_cpus; ++c) {
        const uint64_t* row = base + (uint64_t)c * W;
        const uint64_t v0 = ld_stream_u64(row + lane);
        const uint64_t v1 = has1 ? ld_stream_u64(row + lane + 32) : 0;
        const uint64_t v2 = has2 ? ld_stream_u64(row + lane + 64) : 0;
        s0 += lane == 0 ? (uint64_t)(v0 != E) : v0;
        s1 += v1;
        s2 += v2;
      }
    }
    if (lane == 0) s0 = (s0 == 0) ? (uint64_t)E : 0;  // zero() keeps n_events (:508-512)
    if (a.sums) {
      uint64_t* out = a.sums + i * W;
      if (lane < W) out[lane] = s0;
      if (WIDE && lane + 32 < W) out[lane + 32] = s1;
      if (WIDE && lane + 64 < W) out[lane + 64] = s2;
    }
    if (a.do_scale) {
      const bool ok = __shfl_sync(FULL, s0, 0) != 0;
      const uint64_t te = __shfl_sync(FULL, s0, 1), tr = __shfl_sync(FULL, s0, 2);
      // the scaled count of column j = lane + 32k is event j - 3
      uint64_t c0 = 0, c1 = 0, c2 = 0;
      if (lane >= 3 && lane < W) {
        c0 = ok ? scaled_count(s0, te, tr) : 0;
        if (a.scaled) a.scaled[i * E + lane - 3] = c0;
        if (a.rates) a.rates[i * E + lane - 3] = rate_per_ms(c0, tr);
      }
      if (WIDE && lane + 32 < W) {
        c1 = ok ? scaled_count(s1, te, tr) : 0;
        if (a.scaled) a.scaled[i * E + lane + 29] = c1;
        if (a.rates) a.rates[i * E + lane + 29] = rate_per_ms(c1, tr);
      }
      if (WIDE && lane + 64 < W) {
        c2 = ok ? scaled_count(s2, te, tr) : 0;
        if (a.scaled) a.scaled[i * E + lane + 61] = c2;
        if (a.rates) a.rates[i * E + lane + 61] = rate_per_ms(c2, tr);
      }
      if (a.ratio_out) {
        for (uint32_t r = 0; r < a.n_ratios; ++r) {
          const hbt_ratio q = a.ratios[r];
          const uint32_t jn = 3 + q.num_event, jd = 3 + q.den_event;
          const uint64_t xn0 = __shfl_sync(FULL, c0, jn & 31), xd0 = __shfl_sync(FULL, c0, jd & 31);
          uint64_t xn = xn0, xd = xd0;
          if (WIDE) {
            const uint64_t xn1 = __shfl_sync(FULL, c1, jn & 31), xd1 = __shfl_sync(FULL, c1, jd & 31);
            const uint64_t xn2 = __shfl_sync(FULL, c2, jn & 31), xd2 = __shfl_sync(FULL, c2, jd & 31);
            xn = jn < 32 ? xn0 : (jn < 64 ? xn1 : xn2);
            xd = jd < 32 ? xd0 : (jd < 64 ? xd1 : xd2);
          }
          if (lane == 0) a.ratio_out[i * a.n_ratios + r] = ratio_of(xn, xd);
        }
      }
    }
  }
}

// Per-warp epilogue: s0 holds the sum of column `lane` (column 0: number of failed
// reads); writes sums / scaled / rates / ratios of interval i.
__device__ __forceinline__ void counts_epilogue_narrow(const CountsArgs& a, uint64_t i, uint32_t lane,
                                                       uint64_t s0) {
  constexpr unsigned FULL = 0xffffffffu;
  const uint32_t E = a.n_events, W = 3 + E;
  if (lane == 0) s0 = (s0 == 0) ? (uint64_t)E : 0;  // zero() keeps n_events (:508-512)
  if (a.sums && lane < W) a.sums[i * W + lane] = s0;
  if (!a.do_scale) return;
  const bool ok = __shfl_sync(FULL, s0, 0) != 0;
  const uint64_t te = __shfl_sync(FULL, s0, 1), tr = __shfl_sync(FULL, s0, 2);
  uint64_t c0 = 0;
  if (lane >= 3 && lane < W) {
    c0 = ok ? scaled_count(s0, te, tr) : 0;
    if (a.scaled) a.scaled[i * E + lane - 3] = c0;
    if (a.rates) a.rates[i * E + lane - 3] = rate_per_ms(c0, tr);
  }
  if (a.ratio_out) {
    for (uint32_t r = 0; r < a.n_ratios; ++r) {
      const hbt_ratio q = a.ratios[r];
      const uint64_t xn = __shfl_sync(FULL, c0, (3 + q.num_event) & 31);
      const uint64_t xd = __shfl_sync(FULL, c0, (3 + q.den_event) & 31);
      if (lane == 0) a.ratio_out[i * a.n_ratios + r] = ratio_of(xn, xd);
    }
  }
}

// Streaming variant for W <= 32 and intervals that fit a stage: whole intervals are
// staged in shared memory by 1-D bulk async copies (TMA unit), two stages deep, so
// the bytes in flight do not depend on registers or on how many warps are resident.
// The warps of the CTA reduce one staged interval each -- reading consecutive words
// (conflict-free) with the fixed-column lane mapping of k_counts_reduce -- and run
// the epilogue themselves.  When a stage holds fewer than 8 intervals each interval
// is cut into row slices so that all 8 warps have work; the slices meet in a small
// shared-memory table.  Needs C*W even (16-byte aligned interval starts).
#ifndef HBT_COUNTS_STAGE_KB
#define HBT_COUNTS_STAGE_KB 48
#endif
#ifndef HBT_COUNTS_STAGES
#define HBT_COUNTS_STAGES 2
#endif
constexpr int kCountsStageBytes = HBT_COUNTS_STAGE_KB * 1024;
constexpr int kCountsStages = HBT_COUNTS_STAGES;

static __global__ void __launch_bounds__(256, 2)
    k_counts_reduce_tma(const CountsArgs a, uint32_t per_stage) {
  extern __shared__ __align__(128) uint8_t cs_smem[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(cs_smem + kCountsStages * kCountsStageBytes);
  uint64_t* spart = bar + 8;  // [<= 8 items][32] partial column sums of row slices
  const uint32_t E = a.n_events, W = 3 + E;
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint64_t N = (uint64_t)a.n_cpus * W;
  const uint32_t groups = 32 / W, active = groups * W;
  const bool is0 = lane % W == 0;
  constexpr unsigned FULL = 0xffffffffu;
  // this CTA's contiguous range of intervals, walked in chunks of per_stage intervals
  const uint64_t per_cta = (a.n_intervals + gridDim.x - 1) / gridDim.x;
  const uint64_t i_begin = blockIdx.x * per_cta;
  const uint64_t i_end = min(a.n_intervals, i_begin + per_cta);
  if (i_begin >= i_end) return;
  const uint64_t n_chunks = (i_end - i_begin + per_stage - 1) / per_stage;
  if (tid == 0) {
    for (int s = 0; s < kCountsStages; ++s) mbar_init(&bar[s], 1);
    fence_mbar_init();
  }
  __syncthreads();
  const uint64_t policy = l2_policy_evict_first();
  auto issue = [&](uint64_t k) {
    const int s = (int)(k % kCountsStages);
    const uint64_t i0 = i_begin + k * per_stage;
    const uint32_t n = (uint32_t)min((uint64_t)per_stage, i_end - i0);
    const uint32_t bytes = (uint32_t)(n * N * 8);
    mbar_arrive_expect_tx(&bar[s], bytes);
    bulk_g2s_hint(cs_smem + s * kCountsStageBytes, a.reads + i0 * N, bytes, &bar[s], policy);
  };
  if (tid == 0) {
    for (int k0 = 0; k0 < kCountsStages && (uint64_t)k0 < n_chunks; ++k0) issue(k0);
  }
  uint32_t phase_bits = 0;
  for (uint64_t k = 0; k < n_chunks; ++k) {
    const int s = (int)(k % kCountsStages);
    mbar_wait(&bar[s], (phase_bits >> s) & 1);
    phase_bits ^= 1u << s;
    const uint64_t i0 = i_begin + k * per_stage;
    const uint32_t n = (uint32_t)min((uint64_t)per_stage, i_end - i0);
    const uint32_t sbase = smem_u32(cs_smem + s * kCountsStageBytes);
    uint32_t Q = 1;  // row slices per interval
    while (Q * 2 * n <= 8 && Q * 2 <= a.n_cpus) Q *= 2;
    const uint32_t rows_per = (a.n_cpus + Q - 1) / Q;
    for (uint32_t item = warp; item < n * Q; item += 8) {
      const uint32_t j = item / Q, q = item - j * Q;
      const uint32_t r0 = q * rows_per, r1 = min(a.n_cpus, r0 + rows_per);
      const uint32_t ibase = sbase + (uint32_t)((j * N + (uint64_t)r0 * W) * 8);
      const uint32_t words = (r1 > r0 ? r1 - r0 : 0) * W;
      uint64_t acc = 0;
      if (lane < active) {
#pragma unroll 4
        for (uint32_t f = lane; f < words; f += active) {
          const uint64_t v = lds_u64(ibase + 8 * f);
          acc += is0 ? (uint64_t)(v != E) : v;
        }
      }
      uint64_t s0 = acc;
      for (uint32_t g = 1; g < groups; ++g) {
        const uint64_t o = __shfl_sync(FULL, acc, (lane + g * W) & 31);
        if (lane < W) s0 += o;
      }
      if (Q == 1) counts_epilogue_narrow(a, i0 + j, lane, s0);
      else if (lane < W) spart[item * 32 + lane] = s0;
    }
    if (Q > 1) {
      __syncthreads();
      if (warp < n) {
        uint64_t s0 = 0;
        if (lane < W)
          for (uint32_t q = 0; q < Q; ++q) s0 += spart[(warp * Q + q) * 32 + lane];
        counts_epilogue_narrow(a, i0 + warp, lane, s0);
      }
    }
    fence_proxy_async();
    __syncthreads();  // the stage (and spart) are free again
    if (tid == 0 && k + kCountsStages < n_chunks) issue(k + kCountsStages);
  }
}

// K5 alone, from summed triples (used after the cross-GPU integer reduce).
struct ScaleArgs {
  const uint64_t* sums;  // [I][W]
  uint64_t n_intervals;
  uint32_t n_events;
  const hbt_ratio* ratios;
  uint32_t n_ratios;
  uint64_t* scaled;
  float* rates;
  double* ratio_out;
};

// (static: this header is compiled into more than one translation unit)
static __global__ void __launch_bounds__(256) k_counts_scale(const ScaleArgs a) {
  const uint32_t E = a.n_events;
  const uint32_t W = 3 + E;
  const uint32_t per = E + a.n_ratios;  // work items per interval
  const uint64_t total = a.n_intervals * per;
  for (uint64_t g = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; g < total;
       g += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t i = g / per;
    const uint32_t k = (uint32_t)(g - i * per);
    const uint64_t* s = a.sums + i * W;
    const bool ok = s[0] != 0;
    const uint64_t te = s[1], tr = s[2];
    if (k < E) {
      const uint64_t c = ok ? scaled_count(s[3 + k], te, tr) : 0;
      if (a.scaled) a.scaled[i * E + k] = c;
      if (a.rates) a.rates[i * E + k] = rate_per_ms(c, tr);
    } else if (a.ratio_out) {
      const hbt_ratio r = a.ratios[k - E];
      const uint64_t x = ok ? scaled_count(s[3 + r.num_event], te, tr) : 0;
      const uint64_t y = ok ? scaled_count(s[3 + r.den_event], te, tr) : 0;
      a.ratio_out[i * a.n_ratios + (k - E)] = ratio_of(x, y);
    }
  }
}

// deltas[i][e] = scaled[i][e] - scaled[i-1][e] (u64 wrap): MetricSeries::diff
// (dynolog/src/metric_frame/MetricSeries.h:221-227) over consecutive intervals.
static __global__ void __launch_bounds__(256)
    k_scaled_diff(const uint64_t* scaled, const uint64_t* prev, uint64_t n_intervals, uint32_t E,
                  uint64_t* deltas) {
  const uint64_t total = n_intervals * E;
  for (uint64_t g = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; g < total;
       g += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t p = g >= E ? scaled[g - E] : (prev ? prev[g] : 0);
    deltas[g] = scaled[g] - p;
  }
}

}  // namespace hbt
