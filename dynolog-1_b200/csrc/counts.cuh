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

// One CTA per interval at a time (grid-stride).  Thread t of the first
// `active = floor(BLOCK/W)*W` threads reads words t, t+active, ... of the
// interval's contiguous C*W words, so its column (t % W) is fixed and the
// partial sum lives in a register; consecutive threads read consecutive u64s
// (coalesced 256 B per warp request).
template <int BLOCK>
__global__ void __launch_bounds__(BLOCK) k_counts_reduce(const CountsArgs a) {
  __shared__ uint64_t part[BLOCK];
  __shared__ uint64_t tot[3 + HBT_MAX_EVENTS];
  __shared__ uint64_t sc[HBT_MAX_EVENTS];

  const uint32_t E = a.n_events;
  const uint32_t W = 3 + E;
  const uint32_t active = (BLOCK / W) * W;
  const uint32_t per_col = active / W;
  const uint32_t t = threadIdx.x;
  const uint32_t col = t % W;
  const uint64_t N = (uint64_t)a.n_cpus * W;

  for (uint64_t i = blockIdx.x; i < a.n_intervals; i += gridDim.x) {
    const uint64_t* base = a.reads + i * N;
    uint64_t acc = 0;
    if (t < active) {
      if (col == 0) {
        // word0 = n_events of each T; a mismatch marks a failed read
        // (CpuEventsGroupBase::read, CpuEventsGroup.h:599-614).
#pragma unroll 4
        for (uint64_t f = t; f < N; f += active) acc += (ld_stream_u64(base + f) != E);
      } else {
#pragma unroll 4
        for (uint64_t f = t; f < N; f += active) acc += ld_stream_u64(base + f);
      }
    }
    part[t] = acc;
    __syncthreads();
    if (t < W) {
      uint64_t s = 0;
      for (uint32_t k = 0; k < per_col; ++k) s += part[t + k * W];
      if (t == 0) s = (s == 0) ? (uint64_t)E : 0;  // zero() keeps n_events (:508-512)
      tot[t] = s;
      if (a.sums) a.sums[i * W + t] = s;
    }
    __syncthreads();
    if (a.do_scale) {
      const bool ok = tot[0] != 0;
      if (t < E) {
        const uint64_t c = ok ? scaled_count(tot[3 + t], tot[1], tot[2]) : 0;
        sc[t] = c;
        if (a.scaled) a.scaled[i * E + t] = c;
        if (a.rates) a.rates[i * E + t] = rate_per_ms(c, tot[2]);
      }
      if (a.ratio_out && a.n_ratios) {
        __syncthreads();
        if (t < a.n_ratios) {
          const hbt_ratio r = a.ratios[t];
          a.ratio_out[i * a.n_ratios + t] = ratio_of(sc[r.num_event], sc[r.den_event]);
        }
      }
    }
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
