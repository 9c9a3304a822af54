// Metric tables on the device: dynolog's MetricFrame -- a set of
// MetricSeries rings that share one fixed-interval time axis -- and the window statistics of
// MetricSeriesSlice, for the tables the counter-group kernels produce (scaled counts, rates,
// ratios per interval).
//
// Replaces, for a frame kept in HBM:
//   MetricSeries<T>::addSample / diff / rate / avg / percentile   dynolog/src/metric_frame/MetricSeries.h:148-251
//   MetricFrameVector::addSamples, MetricFrameBase::slice         dynolog/src/metric_frame/MetricFrame.cpp:82-95, MetricFrameBase.cpp:42-52
//   MetricFrameTsUnitFixInterval (time axis, match policies)      dynolog/src/metric_frame/MetricFrameTsUnit.cpp:14-132
//   MetricSeriesSlice<T>::{rate,avg,percentile,diff}              dynolog/src/metric_frame/MetricFrameBase.h:66-91
//
// Layout in HBM: one ring of frame_len 8-byte values per series, series-major
// ([n_series][frame_len]); a series is int64 or double (the reference's MetricSeriesVar).
// The time axis is two host integers (samples held, time of the last one): with a fixed
// interval every other sample time follows from them, exactly as in the reference.
//
// Exactness: integer results are bit-exact.  avg<double> is a left fold in sample order
// (std::accumulate) and is reproduced as such -- one thread walks a window, the parallelism is
// across (window, series) pairs; rate<double> is the reference's expression evaluated in the same
// order; percentile selects the same order statistic nth_element does.
#include <cuda_runtime.h>

#include <cerrno>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include "../../include/hbt_b200.h"
#include "ctx_internal.h"

namespace {

struct FrameView {
  uint64_t* data;           // [n_series][frame_len]
  const uint8_t* is_double; // [n_series]
  uint32_t n_series, frame_len;
  uint32_t head;            // ring index of the oldest sample
  uint32_t size;            // samples held
};

// rows [n_rows][row_stride] (the first n_series words of a row are the samples) appended in row
// order; `skip` leading rows are dropped (they would be overwritten within this very call)
__global__ void __launch_bounds__(256)
    k_frame_append(FrameView f, const uint64_t* __restrict__ rows, uint64_t n_rows, uint32_t row_stride,
                   uint64_t skip, uint32_t tail) {
  const uint64_t total = (n_rows - skip) * f.n_series;
  for (uint64_t g = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; g < total;
       g += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t r = g / f.n_series;
    const uint32_t s = (uint32_t)(g - r * f.n_series);
    const uint32_t slot = (uint32_t)((tail + r) % f.frame_len);
    f.data[(uint64_t)s * f.frame_len + slot] = rows[(skip + r) * row_stride + s];
  }
}

struct QueryArgs {
  FrameView f;
  const hbt_frame_range* ranges;  // device copy
  uint32_t n_windows;
  int64_t period_ns;
  float percentage;
  uint64_t* diff;      // [W][S] raw 8-byte values (int64 or double bits), any may be null
  double* rate_f64;
  int64_t* rate_i64;
  double* avg_f64;
  int64_t* avg_i64;
  uint64_t* pct;
};

__device__ __forceinline__ uint64_t frame_at(const FrameView& f, uint32_t s, uint64_t off) {
  return f.data[(uint64_t)s * f.frame_len + (f.head + off) % f.frame_len];
}

// MetricSeries::rate's private helper (MetricSeries.h:229-249) with the durations as
// steady_clock ticks (nanoseconds): integral R divides the tick counts, floating R goes through
// microsecond counts.  duration == 0 (a one-sample window) divides by zero in the reference;
// here the integral result is 0 and the floating one follows IEEE (inf / nan).
template <class T>
__device__ __forceinline__ int64_t rate_integral(T value, int64_t period, int64_t duration) {
  if (duration == 0) return 0;
  if (period > duration) return (int64_t)(value * (T)(period / duration));
  const int64_t q = duration / period;
  if (q == 0) return 0;
  return (int64_t)(value / (T)q);
}
template <class T>
__device__ __forceinline__ double rate_floating(T value, int64_t period, int64_t duration) {
  // static_cast<R>(value) * double(us(period)) / us(duration): left to right, IEEE double
  return __ddiv_rn(__dmul_rn((double)value, (double)(period / 1000)), (double)(duration / 1000));
}

// one thread per (window, series): diff, rate, avg
__global__ void __launch_bounds__(128) k_frame_stats(const QueryArgs a) {
  const uint64_t total = (uint64_t)a.n_windows * a.f.n_series;
  for (uint64_t g = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; g < total;
       g += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t w = (uint32_t)(g / a.f.n_series), s = (uint32_t)(g - (uint64_t)w * a.f.n_series);
    const hbt_frame_range r = a.ranges[w];
    const uint64_t b = r.start_offset, e = r.end_offset + 1;  // [begin, end) as MetricSeriesSlice passes them
    const int64_t duration = r.end_time_ns - r.start_time_ns;
    const int n = (int)(e - b);  // Iterator::difference_type is int (MetricSeries.h:26)
    const uint64_t first = frame_at(a.f, s, b), last = frame_at(a.f, s, e - 1);
    if (a.f.is_double[s]) {
      const double d = __dsub_rn(__longlong_as_double((long long)last), __longlong_as_double((long long)first));
      if (a.diff) a.diff[g] = (uint64_t)__double_as_longlong(d);
      if (a.rate_f64) a.rate_f64[g] = rate_floating(d, a.period_ns, duration);
      if (a.rate_i64) a.rate_i64[g] = rate_integral(d, a.period_ns, duration);
      if (a.avg_f64 || a.avg_i64) {
        double acc = 0.0;
        for (uint64_t i = b; i < e; ++i) acc = __dadd_rn(acc, __longlong_as_double((long long)frame_at(a.f, s, i)));
        if (a.avg_f64) a.avg_f64[g] = __ddiv_rn(acc, (double)n);
        if (a.avg_i64) a.avg_i64[g] = 0;  // avg<integral> of a double series: not offered
      }
    } else {
      const int64_t d = (int64_t)(last - first);
      if (a.diff) a.diff[g] = (uint64_t)d;
      if (a.rate_f64) a.rate_f64[g] = rate_floating(d, a.period_ns, duration);
      if (a.rate_i64) a.rate_i64[g] = rate_integral(d, a.period_ns, duration);
      if (a.avg_f64 || a.avg_i64) {
        double accd = 0.0;   // std::accumulate(begin, end, double{0}): acc = acc + double(v), in order
        int64_t acci = 0;    // std::accumulate(begin, end, int64_t{0})
        for (uint64_t i = b; i < e; ++i) {
          const int64_t v = (int64_t)frame_at(a.f, s, i);
          accd = __dadd_rn(accd, (double)v);
          acci += v;
        }
        if (a.avg_f64) a.avg_f64[g] = __ddiv_rn(accd, (double)n);
        if (a.avg_i64) a.avg_i64[g] = acci / n;
      }
    }
  }
}

// percentile: the value nth_element leaves at index nth (MetricSeries.h:207-218) is the nth order
// statistic of the window.  One CTA per (window, series): the window is staged in shared memory,
// every thread counts for its candidates how many elements are smaller and how many equal; the
// candidate with less <= nth < less + equal is the answer (all such candidates hold the same value).
__device__ __forceinline__ bool key_less(uint64_t x, uint64_t y, bool dbl) {
  return dbl ? __longlong_as_double((long long)x) < __longlong_as_double((long long)y)
             : (int64_t)x < (int64_t)y;
}
__global__ void __launch_bounds__(256) k_frame_percentile(const QueryArgs a) {
  extern __shared__ uint64_t s_win[];
  const uint64_t total = (uint64_t)a.n_windows * a.f.n_series;
  for (uint64_t g = blockIdx.x; g < total; g += gridDim.x) {
    const uint32_t w = (uint32_t)(g / a.f.n_series), s = (uint32_t)(g - (uint64_t)w * a.f.n_series);
    const hbt_frame_range r = a.ranges[w];
    const uint32_t n = (uint32_t)(r.end_offset + 1 - r.start_offset);
    const bool dbl = a.f.is_double[s] != 0;
    // lround(percentage * (size - 1)): float arithmetic, as written in the reference (:214)
    const uint32_t nth = (uint32_t)lroundf(a.percentage * (float)(n - 1));
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) s_win[i] = frame_at(a.f, s, r.start_offset + i);
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) {
      const uint64_t v = s_win[i];
      uint32_t less = 0, equal = 0;
      for (uint32_t j = 0; j < n; ++j) {
        const uint64_t u = s_win[j];
        less += key_less(u, v, dbl) ? 1u : 0u;
        equal += (!key_less(u, v, dbl) && !key_less(v, u, dbl)) ? 1u : 0u;
      }
      if (less <= nth && nth < less + equal) a.pct[g] = v;  // same value from every writer
    }
  }
}

}  // namespace

struct hbt_frame {
  hbt_ctx* ctx = nullptr;
  uint32_t n_series = 0, frame_len = 0;
  int64_t interval_us = 0;
  uint64_t* d_data = nullptr;
  uint8_t* d_is_double = nullptr;
  std::vector<uint8_t> is_double;
  uint32_t head = 0, size = 0;        // ring state (the same for every series)
  uint64_t sample_count = 0;          // MetricFrameTsUnitFixInterval::sampleCount_ (== size)
  int64_t last_time_ns = 0;           // lastSampleTime_
  // query scratch
  hbt_frame_range* d_ranges = nullptr;
  uint64_t* d_out = nullptr;
  size_t cap_windows = 0;
  uint64_t* d_stage = nullptr;        // host rows staged for hbt_frame_add_samples
  size_t cap_stage = 0;
};

namespace {

#define FCU(call)                                                                         \
  do {                                                                                    \
    cudaError_t e__ = (call);                                                             \
    if (e__ != cudaSuccess)                                                               \
      return hbt_ctx_fail(f->ctx, e__ == cudaErrorMemoryAllocation ? -ENOMEM : -EIO, "%s: %s", #call, \
                          cudaGetErrorString(e__));                                       \
  } while (0)

FrameView view_of(const hbt_frame* f) {
  return FrameView{f->d_data, f->d_is_double, f->n_series, f->frame_len, f->head, f->size};
}

int append_dev(hbt_frame* f, const uint64_t* rows, uint64_t n_rows, uint32_t row_stride, int64_t last_time_ns) {
  if (n_rows == 0) return 0;
  const uint64_t L = f->frame_len;
  const uint64_t skip = n_rows > L ? n_rows - L : 0;  // older rows would be overwritten at once
  const uint64_t n = n_rows - skip;
  // MetricSeries::addSample (MetricSeries.h:148-158): write at the tail, drop the oldest when full
  const uint32_t tail = (uint32_t)((f->head + (uint64_t)f->size + skip) % L);
  const uint64_t total = n * f->n_series;
  int g = (int)((total + 255) / 256);
  const int cap = hbt_ctx_sm_count(f->ctx) * 8;
  if (g > cap) g = cap;
  k_frame_append<<<g, 256, 0, hbt_ctx_stream(f->ctx)>>>(view_of(f), rows, n_rows, row_stride, skip, tail);
  hbt_ctx_count_launches(f->ctx, 1);
  FCU(cudaGetLastError());
  const uint64_t new_total = (uint64_t)f->size + n_rows;
  if (new_total > L) {
    f->head = (uint32_t)((f->head + (new_total - L)) % L);
    f->size = (uint32_t)L;
  } else {
    f->size = (uint32_t)new_total;
  }
  // MetricFrameTsUnitFixInterval::addSample (MetricFrameTsUnit.cpp:25-28), n_rows times
  f->sample_count = f->size;
  f->last_time_ns = last_time_ns;
  return 0;
}

// MetricFrameTsUnitFixInterval::findMatchingOffset (MetricFrameTsUnit.cpp:66-99)
bool match_offset(const hbt_frame* f, int64_t t_ns, int policy, uint64_t* off, int64_t* time_ns) {
  const uint64_t n = f->sample_count;
  if (n == 0) return false;
  const double offset_float =
      (double)(n - 1) + (double)((t_ns - f->last_time_ns) / 1000) / (double)f->interval_us;
  const double top = (double)(n - 1);
  uint64_t o;
  switch (policy) {
    case HBT_MATCH_CLOSEST:
      o = offset_float < 0 ? 0 : offset_float > top ? n - 1 : (uint64_t)std::lround(offset_float);
      break;
    case HBT_MATCH_PREV_CLOSEST:
      if (offset_float < 0) return false;
      o = offset_float > top ? n - 1 : (uint64_t)std::floor(offset_float);
      break;
    case HBT_MATCH_NEXT_CLOSEST:
      if (offset_float > top) return false;
      o = offset_float < 0 ? 0 : (uint64_t)std::ceil(offset_float);
      break;
    default:
      return false;
  }
  *off = o;
  *time_ns = f->last_time_ns - (int64_t)(n - o - 1) * f->interval_us * 1000;  // offsetToTime
  return true;
}

}  // namespace

extern "C" {

int hbt_frame_create(hbt_ctx* ctx, hbt_frame** out, uint32_t n_series, uint32_t frame_len,
                     const uint8_t* series_is_double, int64_t interval_us) {
  if (!ctx || !out) return -EINVAL;
  *out = nullptr;
  if (n_series == 0 || frame_len == 0 || frame_len > HBT_FRAME_MAX_LEN)
    return hbt_ctx_fail(ctx, -EINVAL, "n_series %u / frame_len %u out of range", n_series, frame_len);
  if (interval_us == 0)  // the reference's constructor throws std::invalid_argument (MetricFrameTsUnit.cpp:18-22)
    return hbt_ctx_fail(ctx, -EINVAL, "interval of a fixed-interval time axis cannot be 0");
  hbt_frame* f = new hbt_frame();
  f->ctx = ctx;
  f->n_series = n_series;
  f->frame_len = frame_len;
  f->interval_us = interval_us;
  f->is_double.assign(n_series, 0);
  if (series_is_double)
    for (uint32_t s = 0; s < n_series; ++s) f->is_double[s] = series_is_double[s] ? 1 : 0;
  auto bail = [&](cudaError_t e, const char* what) {
    const int rc = hbt_ctx_fail(ctx, e == cudaErrorMemoryAllocation ? -ENOMEM : -EIO, "%s: %s", what,
                                cudaGetErrorString(e));
    hbt_frame_destroy(f);
    return rc;
  };
  cudaError_t e = cudaSetDevice(hbt_ctx_device(ctx));
  if (e != cudaSuccess) return bail(e, "cudaSetDevice");
  if ((e = cudaMalloc(&f->d_data, (size_t)n_series * frame_len * 8)) != cudaSuccess) return bail(e, "cudaMalloc");
  if ((e = cudaMalloc(&f->d_is_double, n_series)) != cudaSuccess) return bail(e, "cudaMalloc");
  if ((e = cudaMemcpyAsync(f->d_is_double, f->is_double.data(), n_series, cudaMemcpyHostToDevice,
                           hbt_ctx_stream(ctx))) != cudaSuccess)
    return bail(e, "cudaMemcpyAsync");
  if ((e = cudaStreamSynchronize(hbt_ctx_stream(ctx))) != cudaSuccess) return bail(e, "cudaStreamSynchronize");
  *out = f;
  return 0;
}

void hbt_frame_destroy(hbt_frame* f) {
  if (!f) return;
  cudaSetDevice(hbt_ctx_device(f->ctx));
  cudaStreamSynchronize(hbt_ctx_stream(f->ctx));
  if (f->d_data) cudaFree(f->d_data);
  if (f->d_is_double) cudaFree(f->d_is_double);
  if (f->d_ranges) cudaFree(f->d_ranges);
  if (f->d_out) cudaFree(f->d_out);
  if (f->d_stage) cudaFree(f->d_stage);
  delete f;
}

uint64_t hbt_frame_length(const hbt_frame* f) { return f ? f->sample_count : 0; }

int hbt_frame_add_samples_dev(hbt_frame* f, const void* rows, uint64_t n_rows, uint32_t row_stride_words,
                              int64_t last_time_ns) {
  if (!f) return -EINVAL;
  if (n_rows && (!rows || row_stride_words < f->n_series))
    return hbt_ctx_fail(f->ctx, -EINVAL, "rows / row stride");  // addSamples returns false on a width mismatch
  FCU(cudaSetDevice(hbt_ctx_device(f->ctx)));
  return append_dev(f, static_cast<const uint64_t*>(rows), n_rows, row_stride_words, last_time_ns);
}

int hbt_frame_add_samples(hbt_frame* f, const void* rows, uint64_t n_rows, int64_t last_time_ns) {
  if (!f || (n_rows && !rows)) return -EINVAL;
  if (n_rows == 0) return 0;
  FCU(cudaSetDevice(hbt_ctx_device(f->ctx)));
  const size_t words = (size_t)n_rows * f->n_series;
  if (words > f->cap_stage) {
    if (f->d_stage) cudaFree(f->d_stage);
    f->d_stage = nullptr;
    f->cap_stage = 0;
    FCU(cudaMalloc(&f->d_stage, words * 8));
    f->cap_stage = words;
  }
  FCU(cudaMemcpyAsync(f->d_stage, rows, words * 8, cudaMemcpyHostToDevice, hbt_ctx_stream(f->ctx)));
  const int rc = append_dev(f, f->d_stage, n_rows, f->n_series, last_time_ns);
  FCU(cudaStreamSynchronize(hbt_ctx_stream(f->ctx)));  // the staging buffer is reused by the next call
  return rc;
}

int hbt_frame_get_range(const hbt_frame* f, int64_t start_ns, int64_t end_ns, int start_policy, int end_policy,
                        hbt_frame_range* out) {
  if (!f || !out) return -EINVAL;
  hbt_frame_range r{};
  if (!match_offset(f, start_ns, start_policy, &r.start_offset, &r.start_time_ns) ||
      !match_offset(f, end_ns, end_policy, &r.end_offset, &r.end_time_ns))
    return -ENOENT;  // getRange returns std::nullopt (MetricFrameTsUnit.cpp:49-60)
  *out = r;
  return 0;
}

int hbt_frame_query(hbt_frame* f, const hbt_frame_range* ranges, uint32_t n_windows, int64_t period_ns,
                    float percentage, uint64_t* diff, double* rate_f64, int64_t* rate_i64, double* avg_f64,
                    int64_t* avg_i64, uint64_t* pct) {
  if (!f || (n_windows && !ranges)) return -EINVAL;
  if (n_windows == 0) return 0;
  if (period_ns <= 0) return hbt_ctx_fail(f->ctx, -EINVAL, "period must be positive");
  for (uint32_t w = 0; w < n_windows; ++w)
    if (ranges[w].end_offset < ranges[w].start_offset || ranges[w].end_offset >= f->size)
      return hbt_ctx_fail(f->ctx, -EINVAL, "window %u outside the frame", w);
  FCU(cudaSetDevice(hbt_ctx_device(f->ctx)));
  cudaStream_t st = hbt_ctx_stream(f->ctx);
  const size_t per = (size_t)n_windows * f->n_series;
  if (n_windows > f->cap_windows) {
    if (f->d_ranges) cudaFree(f->d_ranges);
    if (f->d_out) cudaFree(f->d_out);
    f->d_ranges = nullptr;
    f->d_out = nullptr;
    f->cap_windows = 0;
    FCU(cudaMalloc(&f->d_ranges, n_windows * sizeof(hbt_frame_range)));
    FCU(cudaMalloc(&f->d_out, 6 * per * 8));
    f->cap_windows = n_windows;
  }
  FCU(cudaMemcpyAsync(f->d_ranges, ranges, n_windows * sizeof(hbt_frame_range), cudaMemcpyHostToDevice, st));
  QueryArgs a{};
  a.f = view_of(f);
  a.ranges = f->d_ranges;
  a.n_windows = n_windows;
  a.period_ns = period_ns;
  a.percentage = percentage;
  uint64_t* o = f->d_out;
  a.diff = diff ? o : nullptr;
  a.rate_f64 = rate_f64 ? reinterpret_cast<double*>(o + per) : nullptr;
  a.rate_i64 = rate_i64 ? reinterpret_cast<int64_t*>(o + 2 * per) : nullptr;
  a.avg_f64 = avg_f64 ? reinterpret_cast<double*>(o + 3 * per) : nullptr;
  a.avg_i64 = avg_i64 ? reinterpret_cast<int64_t*>(o + 4 * per) : nullptr;
  a.pct = pct ? o + 5 * per : nullptr;
  if (a.diff || a.rate_f64 || a.rate_i64 || a.avg_f64 || a.avg_i64) {
    int g = (int)((per + 127) / 128);
    const int cap = hbt_ctx_sm_count(f->ctx) * 16;
    if (g > cap) g = cap;
    k_frame_stats<<<g, 128, 0, st>>>(a);
    hbt_ctx_count_launches(f->ctx, 1);
    FCU(cudaGetLastError());
  }
  if (a.pct) {
    if (percentage < 0.f || percentage > 1.f) return hbt_ctx_fail(f->ctx, -EINVAL, "percentage outside [0, 1]");
    unsigned g = per < (size_t)hbt_ctx_sm_count(f->ctx) * 8 ? (unsigned)per : (unsigned)hbt_ctx_sm_count(f->ctx) * 8;
    const size_t smem = (size_t)f->frame_len * 8;
    if (smem > 48 * 1024)
      FCU(cudaFuncSetAttribute(k_frame_percentile, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_frame_percentile<<<g, 256, smem, st>>>(a);
    hbt_ctx_count_launches(f->ctx, 1);
    FCU(cudaGetLastError());
  }
  auto back = [&](void* dst, size_t slot) {
    return dst ? cudaMemcpyAsync(dst, o + slot * per, per * 8, cudaMemcpyDeviceToHost, st) : cudaSuccess;
  };
  FCU(back(diff, 0));
  FCU(back(rate_f64, 1));
  FCU(back(rate_i64, 2));
  FCU(back(avg_f64, 3));
  FCU(back(avg_i64, 4));
  FCU(back(pct, 5));
  FCU(cudaStreamSynchronize(st));
  return 0;
}

}  // extern "C"
