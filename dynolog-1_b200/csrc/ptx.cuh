// Small sm_100a PTX wrappers used by the decode / counts kernels: streaming
// loads, mbarrier, 1-D bulk async copy (TMA unit, SASS UBLKCP) and the shared-memory 128-bit CAS.
#pragma once
#include <cstdint>

namespace hbt {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// Streaming 8-byte global loads that do not allocate in L1 (data is touched once).
__device__ __forceinline__ uint64_t ld_stream_u64(const uint64_t* p) {
  uint64_t v;
  asm volatile("ld.global.nc.L1::no_allocate.u64 %0, [%1];" : "=l"(v) : "l"(p));
  return v;
}
// ---- mbarrier --------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() {
  // make the init visible to the async (TMA) proxy
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}

// 1-D bulk async copy global -> shared, completion signalled on an mbarrier
// (the linear mode of the TMA unit).  dst/src 16-byte aligned, bytes % 16 == 0.
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes,
                                         uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(smem_u32(smem_dst)),
      "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}

// Same with an L2 eviction-priority hint: ring bytes are read exactly once, so
// they are marked evict_first and leave L2 to the (re-used) histogram table.
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void bulk_g2s_hint(void* smem_dst, const void* gmem_src, uint32_t bytes,
                                              uint64_t* bar, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
      "[%0], [%1], %2, [%3], %4;" ::"r"(smem_u32(smem_dst)),
      "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
      : "memory");
}

// Variants taking shared-space addresses computed once (smem_u32 is a generic->shared
// conversion the compiler otherwise re-emits at every use when registers are tight), and
// pin_u32: an empty volatile asm that makes a value opaque, so that it is kept in its
// register instead of being recomputed (an integer division, a cvta) inside the hot loop.
__device__ __forceinline__ void pin_u32(uint32_t& v) { asm volatile("" : "+r"(v)); }
__device__ __forceinline__ void mbar_arrive_expect_tx_u32(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait_u32(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
  } while (!ok);
}
__device__ __forceinline__ void bulk_g2s_hint_u32(uint32_t smem_dst, const void* gmem_src, uint32_t bytes,
                                                  uint32_t bar, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
      "[%0], [%1], %2, [%3], %4;" ::"r"(smem_dst),
      "l"(gmem_src), "r"(bytes), "r"(bar), "l"(policy)
      : "memory");
}

__device__ __forceinline__ uint64_t lds_u64(uint32_t saddr) {
  uint64_t v;
  asm volatile("ld.shared.u64 %0, [%1];" : "=l"(v) : "r"(saddr) : "memory");
  return v;
}

// Order generic-proxy reads of a stage before the async proxy overwrites it.
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// 128-bit compare-and-swap on shared memory (claims a 16-byte cache slot in one step).
__device__ __forceinline__ void cas128_shared(uint32_t saddr, uint64_t clo, uint64_t chi,
                                              uint64_t nlo, uint64_t nhi, uint64_t& olo,
                                              uint64_t& ohi) {
  asm volatile(
      "{\n\t.reg .b128 c, n, d;\n\t"
      "mov.b128 c, {%3, %4};\n\t"
      "mov.b128 n, {%5, %6};\n\t"
      "atom.shared.relaxed.cta.cas.b128 d, [%2], c, n;\n\t"
      "mov.b128 {%0, %1}, d;\n\t}"
      : "=l"(olo), "=l"(ohi)
      : "r"(saddr), "l"(clo), "l"(chi), "l"(nlo), "l"(nhi)
      : "memory");
}

}  // namespace hbt
