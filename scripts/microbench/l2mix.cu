// Microbenchmark (tuning aid): random CAS.128 on a table of T MB while the same
// CTAs stream a large buffer through shared memory with cp.async.bulk, with and
// without an L2 evict_first hint on the stream.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../dynolog-1_b200/csrc/ptx.cuh"
using namespace hbt;

constexpr int TILE = 16384;
template <int HINT, int ATOM>
__global__ void __launch_bounds__(256) k(const uint8_t* stream, uint64_t stream_bytes, ulonglong2* tab,
                                         uint64_t slots, uint32_t cas_per_tile, uint64_t* sink) {
  extern __shared__ __align__(1024) uint8_t sm[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + 2 * TILE);
  const uint32_t tid = threadIdx.x;
  if (tid == 0) { mbar_init(&bar[0], 1); mbar_init(&bar[1], 1); fence_mbar_init(); }
  __syncthreads();
  const uint64_t pol = l2_policy_evict_first();
  const uint64_t n_tiles = stream_bytes / TILE;
  uint64_t acc = 0;
  uint32_t ph = 0, it = 0;
  for (uint64_t t = blockIdx.x; t < n_tiles; t += gridDim.x, ++it) {
    const int s = it & 1;
    if (tid == 0) {
      mbar_arrive_expect_tx(&bar[s], TILE);
      if (HINT) bulk_g2s_hint(sm + s * TILE, stream + t * TILE, TILE, &bar[s], pol);
      else bulk_g2s(sm + s * TILE, stream + t * TILE, TILE, &bar[s]);
    }
    if (ATOM) {
      for (uint32_t i = tid; i < cas_per_tile; i += 256) {
        uint64_t h = mix64((t * 4096 + i) * 0x9E3779B97F4A7C15ull);
        uint64_t sl = ((h >> 32) * slots) >> 32;
        uint64_t a, b;
        cas128_global(tab + sl, 0, 0, h | 1, 1ull << 32, a, b);
        acc += a ^ b;
      }
    }
    mbar_wait(&bar[s], (ph >> s) & 1);
    ph ^= 1u << s;
    acc += reinterpret_cast<uint64_t*>(sm + s * TILE)[tid];
    fence_proxy_async();
    __syncthreads();
  }
  if (acc == 0x1234567) sink[0] = acc;
}

template <int HINT, int ATOM>
float run(const uint8_t* stream, uint64_t bytes, ulonglong2* tab, uint64_t slots, uint32_t cpt, uint64_t* sink) {
  cudaFuncSetAttribute(k<HINT, ATOM>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * TILE + 64);
  cudaMemset(tab, 0, slots * 16);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<HINT, ATOM><<<148 * 4, 256, 2 * TILE + 64>>>(stream, bytes, tab, slots, cpt, sink);
  cudaMemset(tab, 0, slots * 16);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  k<HINT, ATOM><<<148 * 4, 256, 2 * TILE + 64>>>(stream, bytes, tab, slots, cpt, sink);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  return ms;
}

int main() {
  const uint64_t bytes = 4ull << 30;
  uint8_t* stream; ulonglong2* tab; uint64_t* sink;
  cudaMalloc(&stream, bytes); cudaMemset(stream, 1, bytes);
  cudaMalloc(&tab, 1ull << 31); cudaMalloc(&sink, 8);
  const uint32_t cpt = 410;  // CAS per 16 KB tile ~ one per 40-byte sample
  printf("stream only: nohint %.3f ms, hint %.3f ms\n", run<0, 0>(stream, bytes, tab, 1 << 20, cpt, sink),
         run<1, 0>(stream, bytes, tab, 1 << 20, cpt, sink));
  for (uint64_t mb : {16ull, 64ull, 128ull, 512ull, 2048ull}) {
    uint64_t slots = mb * (1 << 20) / 16;
    float a = run<0, 1>(stream, bytes, tab, slots, cpt, sink);
    float b = run<1, 1>(stream, bytes, tab, slots, cpt, sink);
    double ops = (double)(bytes / TILE) * cpt;
    printf("table %5llu MB: nohint %.3f ms (%.1f Gcas/s), evict_first hint %.3f ms (%.1f Gcas/s)\n",
           (unsigned long long)mb, a, ops / a / 1e6, b, ops / b / 1e6);
  }
  return 0;
}
