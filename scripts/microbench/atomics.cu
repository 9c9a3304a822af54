// Microbenchmark (tuning aid, not product): throughput of the global atomics the
// histogram could use, on tables of different footprints.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o atomics atomics.cu && ./atomics
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../dynolog-1_b200/csrc/ptx.cuh"
using namespace hbt;

template <int OP>
__global__ void k(ulonglong2* tab, uint64_t slots, uint32_t per_thread, uint64_t* sink) {
  uint64_t g = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  uint64_t acc = 0;
  for (uint32_t i = 0; i < per_thread; ++i) {
    uint64_t h = mix64(g * 1315423911ull + i * 2654435761ull);
    uint64_t s = ((h >> 32) * slots) >> 32;
    if (OP == 0) {  // CAS.128, result consumed late
      uint64_t a, b;
      cas128_global(tab + s, 0, 0, h | 1, 1ull << 32, a, b);
      acc += a + b;
    } else if (OP == 1) {  // RED.32
      red_add_u32(reinterpret_cast<uint32_t*>(tab + s) + 3, 1);
    } else if (OP == 2) {  // CAS.64
      acc += atomicCAS(reinterpret_cast<unsigned long long*>(tab + s), 0ull, (unsigned long long)(h | 1));
    } else if (OP == 3) {  // plain 16B load
      ulonglong2 v = ld_relaxed_v2u64(tab + s);
      acc += v.x + v.y;
    } else if (OP == 4) {  // ATOM.ADD.32 with return
      acc += atomicAdd(reinterpret_cast<uint32_t*>(tab + s) + 3, 1u);
    } else if (OP == 5) {  // ATOM.ADD.64 with return
      acc += atomicAdd(reinterpret_cast<unsigned long long*>(tab + s) + 1, 1ull);
    }
  }
  if (acc == 0x1234567) sink[0] = acc;
}

template <int OP>
void run(const char* name, ulonglong2* tab, uint64_t slots, uint64_t* sink) {
  const int blocks = 148 * 8, threads = 256;
  const uint32_t per = 64;
  cudaMemset(tab, 0, slots * 16);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  k<OP><<<blocks, threads>>>(tab, slots, per, sink);
  cudaMemset(tab, 0, slots * 16);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  k<OP><<<blocks, threads>>>(tab, slots, per, sink);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  double ops = (double)blocks * threads * per;
  printf("%-14s slots %9llu (%6.1f MB): %7.3f ms  %7.1f Gops/s\n", name, (unsigned long long)slots,
         slots * 16 / 1e6, ms, ops / ms / 1e6);
}

int main() {
  ulonglong2* tab;
  uint64_t* sink;
  const uint64_t max_slots = 1ull << 27;  // 2 GiB
  cudaMalloc(&tab, max_slots * 16);
  cudaMalloc(&sink, 8);
  for (uint64_t slots : {1ull << 20, 1ull << 22, 1ull << 24, 1ull << 27}) {
    run<0>("cas128", tab, slots, sink);
    run<2>("cas64", tab, slots, sink);
    run<1>("red32", tab, slots, sink);
    run<4>("atom_add32", tab, slots, sink);
    run<5>("atom_add64", tab, slots, sink);
    run<3>("ld128", tab, slots, sink);
  }
  return 0;
}
