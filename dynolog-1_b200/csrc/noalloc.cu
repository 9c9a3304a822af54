// hbt_counts_deltas_static_dev: the one entry point of libhbt_b200.so that never
// allocates.
//
// Who needs it.  dynolog is a monitoring daemon that shares its GPU with the job it
// watches.  Two kinds of caller cannot afford an allocating call in their sampling
// loop: (1) a daemon running beside a training job that has already claimed nearly
// all of HBM -- a cudaMalloc there can fail, or succeed only by making the watched
// job's next allocation fail, and the monitor must never be the reason the job
// dies; (2) a loop with a latency budget -- cudaMalloc / cudaFree serialise against
// all work in flight on the device, so one of them in the middle of a collection
// interval stalls both the monitor and the job.  Every other entry point of this
// library sizes its working buffers on demand (hbt_ctx grows them with cudaMalloc);
// this one keeps its only working buffer in the library image itself.
//
// What it computes.  The per-interval increase of the multiplex-scaled counters of
// a recorded counter-group trace: the same numbers as the `deltas` output of
// hbt_counts_pipeline_dev (group sum over CPUs, GroupReadValues::getCount scaling,
// difference of consecutive intervals).  The scaled counts are the intermediate
// that needs a home between the two kernels; here that home is g_scaled_scratch.
#include <cuda_runtime.h>
#include <atomic>
#include <cerrno>

#include "../../include/hbt_b200.h"
#include "counts.cuh"

using namespace hbt;

// The scratch: scaled[interval][event], at most HBT_STATIC_SCRATCH_WORDS u64.
// 4 MiB -- every GPU that loads the library carries it until the process exits.
static __device__ __align__(256) uint64_t g_scaled_scratch[HBT_STATIC_SCRATCH_WORDS];

namespace {

constexpr int kMaxDevices = 64;
// Device address of g_scaled_scratch on each device, resolved on first use.  In
// host code `g_scaled_scratch` is only a handle for the runtime, not an address a
// kernel could use; cudaGetSymbolAddress turns it into one for the current device.
std::atomic<uint64_t*> g_scratch_addr[kMaxDevices];

int scratch_for_current_device(uint64_t** out) {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return -ENODEV;
  uint64_t* p = g_scratch_addr[dev].load(std::memory_order_acquire);
  if (!p) {
    void* sym = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&sym, g_scaled_scratch);
    if (e != cudaSuccess || !sym) {
      cudaGetLastError();
      return -ENODEV;
    }
    p = static_cast<uint64_t*>(sym);
    g_scratch_addr[dev].store(p, std::memory_order_release);
  }
  *out = p;
  return 0;
}

}  // namespace

extern "C" int hbt_counts_deltas_static_dev(const uint64_t* reads, uint64_t n_intervals,
                                            uint32_t n_cpus, uint32_t n_events, uint64_t* deltas,
                                            void* cuda_stream) {
  if (!reads || !deltas) return -EINVAL;
  if (n_cpus == 0 || n_cpus > HBT_MAX_CPUS || n_events == 0 || n_events > HBT_MAX_EVENTS)
    return -EINVAL;
  if (n_intervals == 0) return 0;
  // the scaled counts of the whole call must fit the fixed scratch
  if (n_intervals > HBT_STATIC_SCRATCH_WORDS / n_events) return -E2BIG;
  uint64_t* scaled = nullptr;
  if (int rc = scratch_for_current_device(&scaled)) return rc;
  cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);

  CountsArgs a{};
  a.reads = reads;
  a.n_intervals = n_intervals;
  a.n_cpus = n_cpus;
  a.n_events = n_events;
  a.ratios = nullptr;
  a.n_ratios = 0;
  a.sums = nullptr;
  a.scaled = scaled;  // every word k_scaled_diff reads below is written here first,
  a.rates = nullptr;  // so nothing left over from an earlier call is ever observed
  a.ratio_out = nullptr;
  a.do_scale = 1;
  int sms = 0, dev = 0;
  cudaGetDevice(&dev);
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0)
    sms = 148;
  uint64_t grid = (uint64_t)sms * 8;  // one warp per interval, 8 warps per CTA
  if (grid > (n_intervals + 7) / 8) grid = (n_intervals + 7) / 8;
  if (3 + n_events <= 32)
    k_counts_reduce<false><<<(unsigned)grid, 256, 0, stream>>>(a);
  else
    k_counts_reduce<true><<<(unsigned)grid, 256, 0, stream>>>(a);
  if (cudaGetLastError() != cudaSuccess) return -EIO;
  const uint64_t total = n_intervals * n_events;
  uint64_t g2 = (total + 255) / 256;
  if (g2 > (uint64_t)sms * 16) g2 = (uint64_t)sms * 16;
  k_scaled_diff<<<(unsigned)g2, 256, 0, stream>>>(scaled, nullptr, n_intervals, n_events, deltas);
  if (cudaGetLastError() != cudaSuccess) return -EIO;
  return 0;
}
