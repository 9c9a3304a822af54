// Glue between hbt's OWN classes and the C-ABI of include/hbt_b200.h: the few lines a dynolog
// maintainer adds (INTEGRATION.md sections 1 and 2), as compiled code.
//
// The functions are templates over the caller's types and include no reference header: they
// need a generator with the member onCpuDataBufferRead(OnRbReadCallback, bool)
// (hbt/src/perf_event/CpuEventsGroup.h:1524-1570; the callback receives RbDataSlices, :44-63)
// and a GroupReadValues<TMode> whose `t` points at the kernel's read_format POD (:371-379).
// oracle/boundary_ref.cpp instantiates them with the reference's real CpuCountSampleGenerator,
// CpuThreadSwitchGenerator and GroupReadValues<mode::Counting>; the mirror classes in
// hbt_gpu.h have the same members and use them too.
#pragma once
#include <linux/perf_event.h>

#include <algorithm>
#include <cerrno>
#include <cstdint>
#include <cstring>

#include "../../include/hbt_b200.h"

namespace hbt_b200 {

// Ring ingest.  Runs the generator's own onCpuDataBufferRead and hands the one or two slices it
// reports to hbt_submit_ring as the unread window of ring `cpu_index`.  `page` is the
// generator's perf_event_mmap_page (mmap_base_): the slices carry pointers and sizes but not
// the positions, and in the wrapped case the reference reports the WHOLE remainder of the ring
// as the second slice (data_size - len, :1561), so its true length comes from data_head.
// consume = false: the tail is advanced later, from hbt_cpu_result::new_tail (writeDataTail).
template <class TCpuGenerator>
int submitUnreadBytes(hbt_ctx* ctx, uint32_t cpu_index, TCpuGenerator& gen,
                      const volatile perf_event_mmap_page* page) {
  const uint64_t head = page->data_head, tail = page->data_tail;  // the values the callee snapshots
  int rc = 0;
  bool called = false;
  gen.onCpuDataBufferRead(
      [&](auto /*cpu*/, const auto& slices) {
        called = true;
        const size_t n1 = slices.first.size;
        size_t n2 = 0;
        const void* p2 = nullptr;
        if (slices.second) {
          n2 = std::min<size_t>(slices.second->size, static_cast<size_t>(head - tail) - n1);
          p2 = slices.second->from;
        }
        rc = hbt_submit_ring(ctx, cpu_index, slices.first.from, n1, p2, n2, tail);
      },
      /*consume=*/false);
  return called ? rc : -ENODATA;
}

// Counter groups.  sums_row is one row [3 + n_events] of hbt_counts_*'s `sums` output: the
// kernel's read_format summed over the CPUs, i.e. exactly what PerCpuBase::read leaves in its
// GroupReadValues (PerCpuBase.h:104-119).  Returns false for a failed read (word0 == 0), for
// which PerCpuCountReader::read returns std::nullopt (PerCpuCountReader.h:112-119).
template <class TReadValues>
bool storeReadValues(TReadValues& rv, const uint64_t* sums_row) {
  const uint64_t n_events = rv.getNumEvents();
  if (sums_row[0] != n_events) return false;
  std::memcpy(rv.t, sums_row, 8 * (3 + n_events));
  return true;
}

}  // namespace hbt_b200
