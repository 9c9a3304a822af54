// TEST INFRASTRUCTURE -- not part of the product path.
//
// ref_harness.cpp: drives the REFERENCE's own headers (included from
// /root/reference, never copied) over synthetic perf_event ring images and
// recorded counter-group traces, and exposes the results through a small C ABI
// (libhbt_ref.so, built into oracle/_ref/ by oracle/Makefile).  It is the
// ground truth the C restatement (oracle/hbt_oracle.c) and the CUDA path are
// pinned against, and the timed CPU baseline of bench.py (--impl reference).
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference
// legs may load it.
//
// How the reference is driven (pattern of
// hbt/src/perf_event/tests/CpuEventsGroupTest.cpp:58-88): a CRTP subclass of
// CpuEventsGroup<Impl, Mode> gets its protected mmap_base_ pointed at a heap
// perf_event_mmap_page followed by the data area, event_fds_[0] set to a valid
// fd so isOpen() holds (CpuEventsGroup.h:581-583), and consume() is called.
// CpuCountSampleGenerator is `final` with protected state, so the headers are
// included under `#define protected public`.

#include <fcntl.h>
#include <linux/perf_event.h>
#include <unistd.h>
#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <iostream>
#include <map>
#include <memory>
#include <optional>
#include <sstream>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include <fmt/format.h>

#define protected public
#include "hbt/src/perf_event/PerCpuCountReader.h"
#include "hbt/src/perf_event/PerCpuCountSampleGenerator.h"
#include "hbt/src/perf_event/PerCpuThreadSwitchGenerator.h"
#include "hbt/src/ringbuffer/Consumer.h"
#undef protected

#include "../include/hbt_b200.h"

namespace hp = facebook::hbt::perf_event;
namespace hb = facebook::hbt;

namespace {

int devnull_fd() {
  static int fd = ::open("/dev/null", O_RDONLY);
  return fd;
}

// A perf_event_mmap_page (4096 B) header as mmap_() finds it
// (CpuEventsGroup.h:1107-1127).  The data area is the caller's ring image, not a
// copy: a header page placed anywhere works because consume() addresses data as
// mmap_base_ + data_offset (CpuEventsGroup.h:1349-1350) and data_offset is a u64
// we choose.
struct FakeMmapAlias {
  alignas(64) uint8_t page[4096];
  perf_event_mmap_page* hdr = nullptr;
  void init(const uint8_t* ring, uint64_t data_size, uint64_t tail, uint64_t head,
            const hbt_tsc_params& tsc) {
    std::memset(page, 0, sizeof(page));
    hdr = reinterpret_cast<perf_event_mmap_page*>(page);
    hdr->data_offset = static_cast<uint64_t>(reinterpret_cast<uintptr_t>(ring)) -
        static_cast<uint64_t>(reinterpret_cast<uintptr_t>(page));  // wraps mod 2^64: fine
    hdr->data_size = data_size;
    hdr->data_head = head;
    hdr->data_tail = tail;
    hdr->time_zero = tsc.time_zero;
    hdr->time_mult = tsc.time_mult;
    hdr->time_shift = tsc.time_shift;
    hdr->cap_user_time = 1;
    hdr->cap_user_time_zero = 1;
  }
};

hp::EventConfs make_confs(uint32_t n_events, uint64_t tsc_mask) {
  hp::EventConfs confs;
  for (uint32_t i = 0; i < n_events; ++i) {
    hp::EventConf c;
    c.id = "ev" + std::to_string(i);
    c.configs = hp::EventConfigs{.type = 0, .config = i};
    if (i < 64 && ((tsc_mask >> i) & 1)) {
      c.transforms.flags = hp::EventValueTransforms::TscToKernelTime;
    }
    confs.push_back(c);
  }
  return confs;
}

template <class G>
void attach(G& g, void* mmap_base, const hbt_tsc_params& tsc) {
  g.mmap_base_ = mmap_base;
  g.event_fds_[0] = devnull_fd();
  g.tsc_rem_flags_ = ((((uint64_t)1) << tsc.time_shift) - 1);  // CpuEventsGroup.h:1146
}

template <class G>
void detach(G& g) {
  // Keep ~CpuEventsGroupBase from close()-ing (munmap, ioctl) fake state.
  g.event_fds_[0] = -1;
  g.mmap_base_ = nullptr;
}

template <class G>
void load_counters(G& g, const uint64_t* c) {
  g.num_record_lost_ = c[HBT_CNT_LOST];
  g.num_record_comm_ = c[HBT_CNT_COMM];
  g.num_record_exit_ = c[HBT_CNT_EXIT];
  g.num_record_throttle_ = c[HBT_CNT_THROTTLE];
  g.num_record_unthrottle_ = c[HBT_CNT_UNTHROTTLE];
  g.num_record_fork_ = c[HBT_CNT_FORK];
  g.num_record_sample_ = c[HBT_CNT_SAMPLE];
  g.num_record_read_ = c[HBT_CNT_READ];
  g.num_record_aux_ = c[HBT_CNT_AUX];
  g.num_record_itrace_start_ = c[HBT_CNT_ITRACE_START];
  g.num_record_switch_cpu_wide_ = c[HBT_CNT_SWITCH_CPU_WIDE];
}

template <class G>
void store_counters(const G& g, uint64_t* c) {
  c[HBT_CNT_LOST] = g.getNumRecordLost();
  c[HBT_CNT_COMM] = g.getNumRecordComm();
  c[HBT_CNT_EXIT] = g.getNumRecordExit();
  c[HBT_CNT_THROTTLE] = g.getNumRecordThrottle();
  c[HBT_CNT_UNTHROTTLE] = g.getNumRecordUnThrottle();
  c[HBT_CNT_FORK] = g.getNumRecordFork();
  c[HBT_CNT_SAMPLE] = g.getNumRecordSample();
  c[HBT_CNT_READ] = g.getNumRecordRead();
  c[HBT_CNT_AUX] = g.getNumRecordAux();
  c[HBT_CNT_ITRACE_START] = g.getNumRecordItraceStart();
  c[HBT_CNT_SWITCH_CPU_WIDE] = g.getNumRecordSwitchCpuWide();
}

// n_records consumed = how far the tail moved is not enough (sizes vary), so a
// counting shim is not available inside consume(); derive it from the return
// value when >= 0, else by re-walking headers from the old to the new tail.
uint64_t count_records_between(const uint8_t* ring, uint64_t data_size, uint64_t from,
                               uint64_t to) {
  uint64_t n = 0;
  const uint64_t mask = data_size - 1;
  while (from < to) {
    auto h = reinterpret_cast<const perf_event_header*>(ring + (from & mask));
    if (h->size == 0) break;
    from += h->size;
    ++n;
  }
  return n;
}

// ---------------------------------------------------------------------------
// Harness-local modes.  consume() is generic over TMode (CpuEventsGroup.h:648-661).
// ---------------------------------------------------------------------------

// Counting-only handlers for the reference's own Sampling mode
// (CpuEventsGroupTest.cpp:58-88 without the logging).
class SamplingCounter final : public hp::CpuEventsGroup<SamplingCounter, hp::mode::Sampling> {
 public:
  using TBase = hp::CpuEventsGroup<SamplingCounter, hp::mode::Sampling>;
  SamplingCounter(hb::CpuId cpu, const hp::EventConfs& confs) : TBase(cpu, -1, -1, confs) {}
  int handleRecordLost(const hp::mode::Sampling::Lost&) noexcept { return 0; }
  int handleRecordRead(const hp::mode::Sampling::Read&) noexcept { return 0; }
  int handleRecordSample(const hp::mode::Sampling::Sample&) noexcept { return 0; }
  void enableImpl() {}
  void disableImpl() {}
};

// Counting-only handlers for ContextSwitch mode (CpuEventsGroupTest.cpp:90-133).
class CtxSwitchCounter final
    : public hp::CpuEventsGroup<CtxSwitchCounter, hp::mode::ContextSwitch> {
 public:
  using TBase = hp::CpuEventsGroup<CtxSwitchCounter, hp::mode::ContextSwitch>;
  using M = hp::mode::ContextSwitch;
  CtxSwitchCounter(hb::CpuId cpu, const hp::EventConfs& confs) : TBase(cpu, -1, -1, confs) {}
  int handleRecordLost(const M::Lost&) noexcept { return 0; }
  int handleRecordComm(const M::Comm_Lower&, size_t, const M::sample_id_t*) noexcept { return 0; }
  int handleRecordExit(const M::Exit&) noexcept { return 0; }
  int handleRecordFork(const M::Fork&) noexcept { return 0; }
  int handleRecordSwitchCpuWide(const M::SwitchCpuWide&) noexcept { return 0; }
  void enableImpl() {}
  void disableImpl() {}
};

// AUX-space mode, counting only.
class AuxCounter final : public hp::CpuEventsGroup<AuxCounter, hp::mode::AUXSpace> {
 public:
  using TBase = hp::CpuEventsGroup<AuxCounter, hp::mode::AUXSpace>;
  using M = hp::mode::AUXSpace;
  AuxCounter(hb::CpuId cpu, const hp::EventConfs& confs) : TBase(cpu, -1, -1, confs) {}
  int handleRecordAux(const M::Aux&) noexcept { return 0; }
  int handleRecordItraceStart(const M::ItraceStart&) noexcept { return 0; }
  void enableImpl() {}
  void disableImpl() {}
};

// Mixed stream of BASELINE config 4: the ContextSwitch records plus the
// Sampling SAMPLE/READ records in one mode.  No such mode exists in the
// reference; consume() itself is the reference's.  The SAMPLE/LOST handlers
// restate CpuCountSampleGenerator's (PerCpuCountSampleGenerator.h:110-202),
// which cannot be reused directly because that class is final and bound to
// mode::Sampling.  PARITY UNPINNED for this mode beyond consume() itself.
struct MixedMode {
  using sample_id_t = hp::mode::ContextSwitch::sample_id_t;
  using Lost = hp::mode::ContextSwitch::Lost;
  using Comm_Lower = hp::mode::ContextSwitch::Comm_Lower;
  using Exit = hp::mode::ContextSwitch::Exit;
  using Fork = hp::mode::ContextSwitch::Fork;
  using Read = hp::mode::Sampling::Read;
  using Sample = hp::mode::Sampling::Sample;
  using Aux = void;
  using ItraceStart = void;
  using SwitchCpuWide = hp::mode::ContextSwitch::SwitchCpuWide;
};

struct DeltaState {
  uint32_t n_events = 0;
  uint64_t tsc_mask = 0;
  uint64_t tstamp = hb::kInvalidTimeStamp;
  uint64_t last[HBT_MAX_EVENTS] = {0};
  uint64_t num_unknown = 0;
  uint64_t* rows = nullptr;
  uint64_t rows_cap = 0;
  uint64_t n_rows = 0;
};

class MixedGroup final : public hp::CpuEventsGroup<MixedGroup, MixedMode> {
 public:
  using TBase = hp::CpuEventsGroup<MixedGroup, MixedMode>;
  MixedGroup(hb::CpuId cpu, const hp::EventConfs& confs) : TBase(cpu, -1, -1, confs) {}
  DeltaState st;
  bool deltas = false;

  int handleRecordLost(const MixedMode::Lost&) noexcept {
    st.tstamp = hb::kInvalidTimeStamp;  // PerCpuCountSampleGenerator.h:122
    return 0;
  }
  int handleRecordComm(const MixedMode::Comm_Lower&, size_t, const MixedMode::sample_id_t*) noexcept {
    return 0;
  }
  int handleRecordExit(const MixedMode::Exit&) noexcept { return 0; }
  int handleRecordFork(const MixedMode::Fork&) noexcept { return 0; }
  int handleRecordSwitchCpuWide(const MixedMode::SwitchCpuWide&) noexcept { return 0; }
  int handleRecordRead(const MixedMode::Read&) noexcept { return 0; }
  ssize_t handleRecordSample(const MixedMode::Sample& r) noexcept {
    if (!deltas) return 0;
    auto xf = [&](uint32_t i, uint64_t v) {
      return ((st.tsc_mask >> i) & 1) ? this->kernelTimeFromTsc(v) : v;
    };
    if (st.tstamp == hb::kInvalidTimeStamp) {  // :135-154
      st.tstamp = r.tstamp;
      for (uint32_t i = 0; i < st.n_events; ++i) st.last[i] = xf(i, r.value[i]);
      ++st.num_unknown;
      return 0;
    }
    if (st.tstamp > r.tstamp) {  // :156-163
      st.tstamp = hb::kInvalidTimeStamp;
      return -EPERM;
    }
    if (st.n_rows >= st.rows_cap) {  // output full: write fails, :191-194
      st.tstamp = hb::kInvalidTimeStamp;
      return -ENOSPC;
    }
    uint64_t* row = st.rows + st.n_rows * (2 + st.n_events);
    row[0] = st.tstamp;
    row[1] = r.tstamp - st.tstamp;  // :165
    for (uint32_t i = 0; i < st.n_events; ++i) {
      uint64_t v = xf(i, r.value[i]);
      row[2 + i] = v - st.last[i];  // :176
      st.last[i] = v;
    }
    ++st.n_rows;
    st.tstamp = r.tstamp;  // :199
    return 0;
  }
  void enableImpl() {}
  void disableImpl() {}
};

// IP sampling (sample_type = IP|TID|TIME|CPU): fixed-size SAMPLE whose size is
// checked like any other non-group record would be, i.e. header.size ==
// sizeof(Sample) through the generic branch of recordSize()
// (CpuEventsGroup.h:328-341).  NO REFERENCE IMPLEMENTATION of the handler
// exists: the histogram below is the build's own
// specification -- PARITY UNPINNED; consume() around it is the reference's.
// consume() finds recordSize() by argument-dependent lookup in the namespace of
// the record type; for harness-local records this overload plays the generic
// `sizeof(TRecord)` branch of the reference's (CpuEventsGroup.h:337-340).
template <class TRecord>
size_t recordSize(const TRecord&) {
  return sizeof(TRecord);
}

template <unsigned kExtraWords>
struct IpMode {
  using sample_id_t = void;
  struct Lost {
    perf_event_header header;
    uint64_t id;
    uint64_t num_lost;
    Lost(const Lost&) = delete;
  };
  using Comm_Lower = void;
  using Exit = void;
  using Fork = void;
  using Read = void;
  struct Sample {
    perf_event_header header;
    uint64_t ip;
    uint32_t pid, tid;
    uint64_t time;
    uint32_t cpu, res;
    uint64_t extra[kExtraWords];
    Sample(const Sample&) = delete;
  };
  using Aux = void;
  using ItraceStart = void;
  using SwitchCpuWide = void;
};

struct IpKey {
  uint64_t ip;
  uint32_t pid;
  bool operator==(const IpKey& o) const { return ip == o.ip && pid == o.pid; }
};
struct IpKeyHash {
  size_t operator()(const IpKey& k) const {
    uint64_t x = k.ip * 0x9E3779B97F4A7C15ull ^ (uint64_t(k.pid) * 0xC2B2AE3D27D4EB4Full);
    x ^= x >> 29;
    return static_cast<size_t>(x);
  }
};
using IpHist = std::unordered_map<IpKey, uint64_t, IpKeyHash>;

template <unsigned kExtraWords>
class IpGroup final : public hp::CpuEventsGroup<IpGroup<kExtraWords>, IpMode<kExtraWords>> {
 public:
  using M = IpMode<kExtraWords>;
  using TBase = hp::CpuEventsGroup<IpGroup<kExtraWords>, M>;
  IpGroup(hb::CpuId cpu, const hp::EventConfs& confs) : TBase(cpu, -1, -1, confs) {}
  IpHist* hist = nullptr;
  int handleRecordLost(const typename M::Lost&) noexcept { return 0; }
  int handleRecordSample(const typename M::Sample& r) noexcept {
    if (hist) ++(*hist)[IpKey{r.ip, r.pid}];
    return 0;
  }
  void enableImpl() {}
  void disableImpl() {}
};

template <class G>
int64_t run_consume(G& g, uint32_t max_records) {
  return static_cast<int64_t>(g.consume(max_records));
}

template <class G>
void finish_result(G& g, const uint8_t* ring, uint64_t data_size, uint64_t old_tail,
                   perf_event_mmap_page* hdr, int64_t ret, hbt_cpu_result* res) {
  res->ret = ret;
  res->new_tail = hdr->data_tail;
  res->n_records = ret >= 0 ? static_cast<uint64_t>(ret)
                            : count_records_between(ring, data_size, old_tail, hdr->data_tail);
  store_counters(g, res->counters);
}

}  // namespace

extern "C" {

int hbtref_abi_version(void) { return HBT_B200_ABI_VERSION; }

// consume() with count-only handlers.  mode: 0 Sampling, 1 ContextSwitch,
// 2 AUXSpace, 3 Mixed (harness-local).  res->counters are in/out.
int hbtref_consume_count(const uint8_t* ring, uint64_t data_size, uint64_t tail, uint64_t head,
                         int mode, uint32_t max_records, hbt_cpu_result* res) {
  hbt_tsc_params tsc{0, 1, 0, 0};
  FakeMmapAlias mm;
  mm.init(ring, data_size, tail, head, tsc);
  auto confs = make_confs(1, 0);
  auto go = [&](auto& g) {
    attach(g, mm.page, tsc);
    load_counters(g, res->counters);
    int64_t ret = run_consume(g, max_records);
    finish_result(g, ring, data_size, tail, mm.hdr, ret, res);
    detach(g);
  };
  switch (mode) {
    case 0: { SamplingCounter g(0, confs); go(g); } break;
    case 1: { CtxSwitchCounter g(0, confs); go(g); } break;
    case 2: { AuxCounter g(0, confs); go(g); } break;
    case 3: { MixedGroup g(0, confs); go(g); } break;
    default: return -EINVAL;
  }
  return 0;
}

// The REAL CpuCountSampleGenerator (PerCpuCountSampleGenerator.h:27-231) over a
// ring image.  carry_tstamp / carry_last are count_.tstamp / last_[] in and out
// (the caller seeds what enableImpl() would, :91-106); rows are drained from
// the generator's own output ring with ringbuffer::Consumer.
int hbtref_sample_deltas(const uint8_t* ring, uint64_t data_size, uint64_t tail, uint64_t head,
                         uint32_t n_events, uint64_t tsc_mask, const hbt_tsc_params* tsc,
                         uint32_t max_records, uint64_t* carry_tstamp, uint64_t* carry_last,
                         uint64_t* rows, uint64_t rows_cap, hbt_cpu_result* res) {
  if (n_events == 0 || n_events > HBT_MAX_EVENTS) return -EINVAL;
  FakeMmapAlias mm;
  mm.init(ring, data_size, tail, head, *tsc);
  auto confs = make_confs(n_events, tsc_mask);
  std::vector<std::string> names;
  for (auto& c : confs) names.push_back(c.id);
  // Output ring sized for every possible row of the window (+1: the ring keeps
  // one byte free); no drop-if-full.
  const uint64_t max_rows = (head - tail) / (40 + 8ull * n_events) + 2;
  hp::CpuCountSampleGenerator g(names, 0, -1, confs, max_rows + 1, 0);
  attach(g, mm.page, *tsc);
  load_counters(g, res->counters);
  g.count_.tstamp = *carry_tstamp;
  for (uint32_t i = 0; i < n_events; ++i) g.last_[i] = carry_last[i];
  g.num_unknown_count_interval_ = static_cast<uint32_t>(res->num_unknown_count_interval);

  int64_t ret = run_consume(g, max_records);
  finish_result(g, ring, data_size, tail, mm.hdr, ret, res);

  const size_t row_bytes = g.kWriteByteSize;
  facebook::hbt::ringbuffer::Consumer<hp::CpuCountSampleGenerator::RbExtraData> cons(g.getOutput());
  uint64_t n = 0;
  std::vector<uint64_t> row(2 + n_events);
  int rc = 0;
  while (true) {
    auto r = cons.copy(row_bytes, row.data());
    if (r < 0) break;
    if (n >= rows_cap) { rc = -ENOSPC; break; }
    std::memcpy(rows + n * (2 + n_events), row.data(), row_bytes);
    ++n;
  }
  res->n_rows = n;
  res->num_unknown_count_interval = g.num_unknown_count_interval_;
  res->carry_tstamp = g.count_.tstamp;
  *carry_tstamp = g.count_.tstamp;
  for (uint32_t i = 0; i < n_events; ++i) carry_last[i] = g.last_[i];
  detach(g);
  return rc;
}

// Mixed-mode consume with the restated delta handlers (see MixedGroup).
int hbtref_mixed_deltas(const uint8_t* ring, uint64_t data_size, uint64_t tail, uint64_t head,
                        uint32_t n_events, uint64_t tsc_mask, const hbt_tsc_params* tsc,
                        uint32_t max_records, uint64_t* carry_tstamp, uint64_t* carry_last,
                        uint64_t* rows, uint64_t rows_cap, hbt_cpu_result* res) {
  if (n_events == 0 || n_events > HBT_MAX_EVENTS) return -EINVAL;
  FakeMmapAlias mm;
  mm.init(ring, data_size, tail, head, *tsc);
  auto confs = make_confs(n_events, tsc_mask);
  MixedGroup g(0, confs);
  attach(g, mm.page, *tsc);
  load_counters(g, res->counters);
  g.deltas = true;
  g.st.n_events = n_events;
  g.st.tsc_mask = tsc_mask;
  g.st.tstamp = *carry_tstamp;
  for (uint32_t i = 0; i < n_events; ++i) g.st.last[i] = carry_last[i];
  g.st.num_unknown = res->num_unknown_count_interval;
  g.st.rows = rows;
  g.st.rows_cap = rows_cap;
  int64_t ret = run_consume(g, max_records);
  finish_result(g, ring, data_size, tail, mm.hdr, ret, res);
  res->n_rows = g.st.n_rows;
  res->num_unknown_count_interval = g.st.num_unknown;
  res->carry_tstamp = g.st.tstamp;
  *carry_tstamp = g.st.tstamp;
  for (uint32_t i = 0; i < n_events; ++i) carry_last[i] = g.st.last[i];
  detach(g);
  return 0;
}

// IP-sample histogram for one ring; entries appended to an opaque map.
void* hbtref_hist_new(void) { return new std::vector<IpHist>(); }
void hbtref_hist_free(void* h) { delete static_cast<std::vector<IpHist>*>(h); }

static int ip_consume_one(const uint8_t* ring, uint64_t data_size, uint64_t tail, uint64_t head,
                          uint32_t ip_sample_bytes, uint32_t max_records, IpHist* hist,
                          hbt_cpu_result* res) {
  hbt_tsc_params tsc{0, 1, 0, 0};
  FakeMmapAlias mm;
  mm.init(ring, data_size, tail, head, tsc);
  auto confs = make_confs(1, 0);
  auto go = [&](auto& g) {
    attach(g, mm.page, tsc);
    load_counters(g, res->counters);
    g.hist = hist;
    int64_t ret = run_consume(g, max_records);
    finish_result(g, ring, data_size, tail, mm.hdr, ret, res);
    detach(g);
  };
  switch (ip_sample_bytes) {
    case 40: { IpGroup<0> g(0, confs); go(g); } break;
    case 48: { IpGroup<1> g(0, confs); go(g); } break;
    case 64: { IpGroup<3> g(0, confs); go(g); } break;
    case 104: { IpGroup<8> g(0, confs); go(g); } break;
    case 128: { IpGroup<11> g(0, confs); go(g); } break;
    default: return -EINVAL;
  }
  return 0;
}

// Decode n_cpus rings laid out back to back (ring c at rings + c*data_size) with
// `threads` host threads; per-CPU histograms keyed (pid, ip).  Returns seconds
// spent in decode+aggregate (steady_clock), or a negative errno.
double hbtref_ip_hist(void* hist_handle, const uint8_t* rings, uint64_t data_size,
                      const uint64_t* tails, const uint64_t* heads, uint32_t n_cpus,
                      uint32_t ip_sample_bytes, uint32_t max_records, uint32_t threads,
                      hbt_cpu_result* res) {
  auto* hists = static_cast<std::vector<IpHist>*>(hist_handle);
  hists->assign(n_cpus, IpHist());
  if (threads == 0) threads = 1;
  std::atomic<uint32_t> next{0};
  std::atomic<int> err{0};
  auto t0 = std::chrono::steady_clock::now();
  auto work = [&]() {
    for (;;) {
      uint32_t c = next.fetch_add(1);
      if (c >= n_cpus) return;
      int rc = ip_consume_one(rings + uint64_t(c) * data_size, data_size, tails[c], heads[c],
                              ip_sample_bytes, max_records, &(*hists)[c], &res[c]);
      if (rc < 0) err = rc;
    }
  };
  std::vector<std::thread> ts;
  for (uint32_t t = 1; t < threads; ++t) ts.emplace_back(work);
  work();
  for (auto& t : ts) t.join();
  auto t1 = std::chrono::steady_clock::now();
  if (err.load() < 0) return static_cast<double>(err.load());
  return std::chrono::duration<double>(t1 - t0).count();
}

uint64_t hbtref_hist_size(void* hist_handle) {
  auto* hists = static_cast<std::vector<IpHist>*>(hist_handle);
  uint64_t n = 0;
  for (auto& h : *hists) n += h.size();
  return n;
}

uint64_t hbtref_hist_fetch(void* hist_handle, hbt_hist_entry* out, uint64_t cap) {
  auto* hists = static_cast<std::vector<IpHist>*>(hist_handle);
  uint64_t n = 0;
  for (uint32_t c = 0; c < hists->size(); ++c) {
    for (auto& kv : (*hists)[c]) {
      if (n >= cap) return n;
      out[n++] = hbt_hist_entry{kv.first.ip, kv.first.pid, c, kv.second};
    }
  }
  return n;
}

// Multi-threaded timed run of the real CpuCountSampleGenerator over n_cpus rings
// (seeded invalid, rows kept in the generator's own ring and discarded).
double hbtref_sample_deltas_timed(const uint8_t* rings, uint64_t data_size, const uint64_t* tails,
                                  const uint64_t* heads, uint32_t n_cpus, uint32_t n_events,
                                  uint32_t threads, uint64_t* total_records, uint64_t* total_rows) {
  if (threads == 0) threads = 1;
  hbt_tsc_params tsc{0, 1, 0, 0};
  std::atomic<uint32_t> next{0};
  std::atomic<uint64_t> recs{0}, rows{0};
  // Generators (and their output rings) are built outside the timed region.
  auto confs = make_confs(n_events, 0);
  std::vector<std::string> names;
  for (auto& c : confs) names.push_back(c.id);
  std::vector<std::unique_ptr<hp::CpuCountSampleGenerator>> gens(n_cpus);
  std::vector<std::unique_ptr<FakeMmapAlias>> mms(n_cpus);
  for (uint32_t c = 0; c < n_cpus; ++c) {
    const uint64_t max_rows = (heads[c] - tails[c]) / (40 + 8ull * n_events) + 2;
    gens[c] = std::make_unique<hp::CpuCountSampleGenerator>(names, c % HBT_MAX_CPUS, -1, confs,
                                                            max_rows + 1, 0);
    mms[c] = std::make_unique<FakeMmapAlias>();
    mms[c]->init(rings + uint64_t(c) * data_size, data_size, tails[c], heads[c], tsc);
    attach(*gens[c], mms[c]->page, tsc);
    gens[c]->count_.tstamp = hb::kInvalidTimeStamp;
  }
  auto t0 = std::chrono::steady_clock::now();
  auto work = [&]() {
    for (;;) {
      uint32_t c = next.fetch_add(1);
      if (c >= n_cpus) return;
      auto& g = *gens[c];
      g.consume(0xFFFFFFFFu);
      recs += g.getNumRecordSample() + g.getNumRecordRead();
      rows += g.getNumEntries();
    }
  };
  std::vector<std::thread> ts;
  for (uint32_t t = 1; t < threads; ++t) ts.emplace_back(work);
  work();
  for (auto& t : ts) t.join();
  auto t1 = std::chrono::steady_clock::now();
  for (auto& g : gens) detach(*g);
  *total_records = recs.load();
  *total_rows = rows.load();
  return std::chrono::duration<double>(t1 - t0).count();
}

// The generator's output ring as a row sink: n_rows rows {tstamp,duration,deltas[n]} go
// through the REAL ringbuffer::Producer of a RingBuffer sized like the generator sizes its
// own (rb_min_num_entries * kWriteByteSize, PerCpuCountSampleGenerator.h:69-71) with the
// write / dropN / retry sequence of handleRecordSample (:182-195); what is left is drained
// with the real ringbuffer::Consumer.  refused = rows whose write failed.
int hbtref_row_sink(const uint64_t* rows, uint64_t n_rows, uint32_t n_events, uint64_t rb_min_num_entries,
                    uint64_t num_counts_drop_if_full, uint64_t* out, uint64_t out_cap, uint64_t* n_out,
                    uint64_t* refused) {
  namespace rb = facebook::hbt::ringbuffer;
  using Extra = hp::CpuCountSampleGenerator::RbExtraData;
  const size_t row_bytes = sizeof(uint64_t) * (2 + n_events);
  const size_t drop_bytes = num_counts_drop_if_full * row_bytes;
  rb::Producer<Extra> prod(std::make_shared<rb::RingBuffer<Extra>>(rb_min_num_entries * row_bytes));
  *refused = 0;
  for (uint64_t r = 0; r < n_rows; ++r) {
    auto ret = prod.write(rows + r * (2 + n_events), row_bytes);
    if (drop_bytes > 0 && ret == -ENOSPC) {
      if (prod.dropN(drop_bytes) < 0) return -EIO;
      ret = prod.write(rows + r * (2 + n_events), row_bytes);
    }
    if (ret < 0) ++*refused;
  }
  rb::Consumer<Extra> cons(prod.getRingBuffer());
  uint64_t n = 0;
  std::vector<uint64_t> row(2 + n_events);
  while (cons.copy(row_bytes, row.data()) >= 0) {
    if (n >= out_cap) return -ENOSPC;
    std::memcpy(out + n * (2 + n_events), row.data(), row_bytes);
    ++n;
  }
  *n_out = n;
  return 0;
}

// The REAL CpuThreadSwitchGenerator (PerCpuThreadSwitchGenerator.h:168-454) over n_cpus
// ring images, CPUs in ascending order sharing one fresh PerThreadStats, exactly like
// PerCpuThreadSwitchGenerator builds them (:466-481).  Events are drained from each
// generator's own output ring; their tags are the reference's virtual ids.
// last_tstamp / in_write_errors are last_tstamp_ / in_write_errors_ in and out.
// The stream must be one the reference accepts: its handlers HBT_THROW_ASSERT inside
// noexcept functions (std::terminate) on a second FORK of a live thread, for example.
int hbtref_switch_events(const uint8_t* rings, uint64_t data_size, const uint64_t* tails,
                         const uint64_t* heads, uint32_t n_cpus, uint32_t cpu_base, uint32_t tid_level,
                         uint32_t max_records, uint64_t* last_tstamp, uint8_t* in_write_errors,
                         hbt_switch_event* events, uint64_t events_cap, uint64_t* offsets,
                         hbt_cpu_result* res) {
  static_assert(sizeof(hbt_switch_event) == sizeof(facebook::hbt::tagstack::Event));
  namespace rb = facebook::hbt::ringbuffer;
  auto stats = std::make_shared<hp::PerThreadStats>();
  hbt_tsc_params tsc{0, 1, 0, 0};
  uint64_t n = 0;
  for (uint32_t c = 0; c < n_cpus; ++c) {
    offsets[c] = n;
    const uint8_t* ring = rings + uint64_t(c) * data_size;
    FakeMmapAlias mm;
    mm.init(ring, data_size, tails[c], heads[c], tsc);
    const uint64_t max_events = 2 * ((heads[c] - tails[c]) / 32 + 2);
    hp::CpuThreadSwitchGenerator g(cpu_base + c, -1, static_cast<facebook::hbt::tagstack::Level>(tid_level),
                                   stats, max_events + 1);
    attach(g, mm.page, tsc);
    load_counters(g, res[c].counters);
    g.last_tstamp_ = last_tstamp[c];
    g.in_write_errors_ = in_write_errors[c] != 0;
    int64_t ret = run_consume(g, max_records);
    finish_result(g, ring, data_size, tails[c], mm.hdr, ret, &res[c]);
    rb::Consumer<hp::CpuThreadSwitchGenerator::RbExtraData> cons(g.getOutput());
    facebook::hbt::tagstack::Event ev;
    uint64_t k = 0;
    while (cons.copy(sizeof(ev), &ev) >= 0) {
      if (n >= events_cap) { detach(g); return -ENOSPC; }
      std::memcpy(&events[n++], &ev, sizeof(ev));
      ++k;
    }
    res[c].n_rows = k;
    res[c].carry_tstamp = g.last_tstamp_;
    res[c].num_unknown_count_interval = g.in_write_errors_ ? 1 : 0;
    last_tstamp[c] = g.last_tstamp_;
    in_write_errors[c] = g.in_write_errors_ ? 1 : 0;
    detach(g);
  }
  offsets[n_cpus] = n;
  return 0;
}

// ---------------------------------------------------------------------------
// Counter groups: the real PerCpuBase<T>::read (PerCpuBase.h:104-119) over a
// fake per-CPU reader that replays a recorded trace, then the real
// GroupReadValues::getCount (CpuEventsGroup.h:449-459).
// ---------------------------------------------------------------------------
}  // extern "C"

namespace {

struct ReplayCpuReader {
  using TMode = hp::mode::Counting;
  const uint64_t* cur = nullptr;  // the T this CPU's next read() returns
  uint32_t n_events = 0;
  bool open = true;
  bool isOpen() const { return open; }
  bool isEnabled() const { return true; }
  void close() {}
  void enable() {}
  void disable() {}
  // CpuEventsGroupBase::read (CpuEventsGroup.h:596-615): the kernel fills rv.t;
  // a short read is a failure.  A recorded T whose n_events differs stands for
  // a failed read.
  bool read(hp::GroupReadValues<TMode>& rv) const {
    if (cur[0] != n_events) return false;
    std::memcpy(rv.t, cur, 8 * (3 + n_events));
    return true;
  }
};

struct ReplayPerCpu : public hp::PerCpuBase<ReplayCpuReader> {
  using TBase = hp::PerCpuBase<ReplayCpuReader>;
  ReplayPerCpu(const hb::CpuSet& cpus, uint32_t n_events) : TBase(cpus, nullptr) {
    for_each_cpu(cpu, cpus) {
      cpu_generators_[cpu] = std::make_shared<ReplayCpuReader>();
      cpu_generators_[cpu]->n_events = n_events;
    }
  }
};

}  // namespace

extern "C" {

// reads [I][C][3+E] -> sums [I][3+E] (word0 = E, or 0 when the read failed),
// scaled [I][E], rates [I][E] (PerfMonitor.cpp:56-57 expression), ratios [I][R],
// deltas [I][E] (u64 difference of consecutive scaled values; interval 0
// against 0).  Any output may be NULL.  Intervals [i0, i1) only (threads split
// the range); the caller handles the delta seam by passing i0 > 0.
static void counts_range(const uint64_t* reads, uint64_t i0, uint64_t i1, uint32_t C, uint32_t E,
                         const hbt_ratio* ratios, uint32_t R, uint64_t* sums, uint64_t* scaled,
                         float* rates, double* ratio_out, uint64_t* deltas) {
  std::vector<unsigned> ids(C);
  for (uint32_t c = 0; c < C; ++c) ids[c] = c;
  hb::CpuSet cpus(hb::CpuSet::make_cpu_set_t(ids), C - 1, C - 1);
  ReplayPerCpu per_cpu(cpus, E);
  hp::GroupReadValues<hp::mode::Counting> rv(E);
  const uint64_t W = 3 + E;
  std::vector<uint64_t> prev(E, 0), cur(E, 0);
  bool have_prev = false;
  // deltas at i0 need scaled[i0-1]: recompute it.
  uint64_t start = (deltas && i0 > 0) ? i0 - 1 : i0;
  for (uint64_t i = start; i < i1; ++i) {
    for (uint32_t c = 0; c < C; ++c) per_cpu.cpu_generators_[c]->cur = reads + (i * C + c) * W;
    bool ok = per_cpu.read(rv);
    const bool emit = i >= i0;
    if (emit && sums) {
      std::memcpy(sums + i * W, rv.t, 8 * W);
      if (!ok) sums[i * W] = 0;
    }
    const uint64_t time = rv.getTimeRunning();  // PerfMonitor.cpp:52
    for (uint32_t e = 0; e < E; ++e) {
      uint64_t count = ok ? rv.getCount(e) : 0;  // PerfMonitor.cpp:55
      cur[e] = count;
      if (!emit) continue;
      if (scaled) scaled[i * E + e] = count;
      if (rates) {
        // PerfMonitor.cpp:56-57 / :62-63
        rates[i * E + e] = static_cast<float>(static_cast<double>(count) * 1e3 /
                                              static_cast<double>(time));
      }
      if (deltas) deltas[i * E + e] = count - ((have_prev || i > 0) ? prev[e] : 0);
    }
    if (emit && ratio_out) {
      for (uint32_t r = 0; r < R; ++r) {
        ratio_out[i * R + r] = static_cast<double>(cur[ratios[r].num_event]) /
            static_cast<double>(cur[ratios[r].den_event]);
      }
    }
    prev = cur;
    have_prev = true;
  }
}

double hbtref_counts(const uint64_t* reads, uint64_t n_intervals, uint32_t n_cpus,
                     uint32_t n_events, const hbt_ratio* ratios, uint32_t n_ratios,
                     uint64_t* sums, uint64_t* scaled, float* rates, double* ratio_out,
                     uint64_t* deltas, uint32_t threads) {
  if (n_cpus == 0 || n_cpus > HBT_MAX_CPUS || n_events == 0 || n_events > HBT_MAX_EVENTS)
    return -EINVAL;
  if (threads == 0) threads = 1;
  if (threads > n_intervals) threads = n_intervals ? static_cast<uint32_t>(n_intervals) : 1;
  auto t0 = std::chrono::steady_clock::now();
  std::vector<std::thread> ts;
  for (uint32_t t = 0; t < threads; ++t) {
    uint64_t a = n_intervals * t / threads, b = n_intervals * (t + 1) / threads;
    auto fn = [=]() {
      counts_range(reads, a, b, n_cpus, n_events, ratios, n_ratios, sums, scaled, rates,
                   ratio_out, deltas);
    };
    if (t + 1 == threads) fn(); else ts.emplace_back(fn);
  }
  for (auto& t : ts) t.join();
  auto t1 = std::chrono::steady_clock::now();
  return std::chrono::duration<double>(t1 - t0).count();
}

// kernelTimeFromTsc through a real generator object (CpuEventsGroup.h:1577-1591).
uint64_t hbtref_kernel_time_from_tsc(uint64_t tsc_cycles, const hbt_tsc_params* tsc) {
  uint8_t dummy[64] = {0};
  FakeMmapAlias mm;
  mm.init(dummy, 64, 0, 0, *tsc);
  auto confs = make_confs(1, 0);
  SamplingCounter g(0, confs);
  attach(g, mm.page, *tsc);
  uint64_t v = g.kernelTimeFromTsc(tsc_cycles);
  detach(g);
  return v;
}

}  // extern "C"

// ---------------------------------------------------------------------------
// Metric tables: the REAL dynolog::MetricFrameVector / MetricSeries / MetricFrameTsUnitFixInterval
// (dynolog/src/metric_frame/*, compiled from the reference's own .cpp files by oracle/Makefile).
// rows [n_rows][S] are added oldest first, row r at t_last - (n_rows-1-r) * interval; each query
// slices the frame like MetricFrameBase::slice and evaluates MetricSeriesSlice's statistics.
// ---------------------------------------------------------------------------
#include "dynolog/src/metric_frame/MetricFrame.h"
#include "dynolog/src/metric_frame/MetricFrameTsUnit.h"

extern "C" int hbtref_frame_query(const uint64_t* rows, uint64_t n_rows, uint32_t S, const uint8_t* is_double,
                                  uint32_t frame_len, int64_t interval_us, int64_t t_last_ns,
                                  const int64_t* q_start, const int64_t* q_end, const int32_t* pol_s,
                                  const int32_t* pol_e, uint32_t Q, int64_t period_ns, float percentage,
                                  hbt_frame_range* ranges_out, uint8_t* ok, uint64_t* diff, double* rate_f64,
                                  int64_t* rate_i64, double* avg_f64, int64_t* avg_i64, uint64_t* pct,
                                  uint64_t* length_out) {
  namespace dl = facebook::dynolog;
  using ns = std::chrono::nanoseconds;
  auto ts = std::make_shared<dl::MetricFrameTsUnitFixInterval>(std::chrono::microseconds(interval_us), frame_len);
  dl::VectorSeriesDefList defs;
  for (uint32_t s = 0; s < S; ++s) {
    if (is_double[s])
      defs.push_back(std::make_shared<dl::MetricSeries<double>>(frame_len, "s" + std::to_string(s), ""));
    else
      defs.push_back(std::make_shared<dl::MetricSeries<int64_t>>(frame_len, "s" + std::to_string(s), ""));
  }
  dl::MetricFrameVector frame(defs, "frame", "", ts);
  for (uint64_t r = 0; r < n_rows; ++r) {
    dl::VectorSamplesT samples;
    for (uint32_t s = 0; s < S; ++s) {
      const uint64_t raw = rows[r * S + s];
      if (is_double[s]) {
        double d;
        std::memcpy(&d, &raw, 8);
        samples.emplace_back(d);
      } else {
        samples.emplace_back(static_cast<int64_t>(raw));
      }
    }
    const int64_t t = t_last_ns - static_cast<int64_t>(n_rows - 1 - r) * interval_us * 1000;
    if (!frame.addSamples(samples, dl::TimePoint(ns(t)))) return -EINVAL;
  }
  *length_out = frame.length();
  const auto period = std::chrono::steady_clock::duration(ns(period_ns));
  for (uint32_t q = 0; q < Q; ++q) {
    const auto a = dl::TimePoint(ns(q_start[q])), b = dl::TimePoint(ns(q_end[q]));
    const auto ps = static_cast<dl::MATCH_POLICY>(pol_s[q]), pe = static_cast<dl::MATCH_POLICY>(pol_e[q]);
    auto range = ts->getRange(a, b, ps, pe);
    auto slice = frame.slice(a, b, ps, pe);
    ok[q] = range.has_value() && slice.has_value() && range->end.offset >= range->start.offset;
    if (!ok[q]) continue;
    ranges_out[q].start_offset = range->start.offset;
    ranges_out[q].end_offset = range->end.offset;
    ranges_out[q].start_time_ns = std::chrono::duration_cast<ns>(range->start.time.time_since_epoch()).count();
    ranges_out[q].end_time_ns = std::chrono::duration_cast<ns>(range->end.time.time_since_epoch()).count();
    const bool zero_duration = range->start.time == range->end.time;  // integral rate would divide by zero
    for (uint32_t s = 0; s < S; ++s) {
      const size_t g = static_cast<size_t>(q) * S + s;
      if (is_double[s]) {
        auto sl = slice->series<double>(static_cast<int>(s));
        if (!sl) return -EIO;
        const double d = sl->diff(), p = sl->percentile(percentage);
        std::memcpy(&diff[g], &d, 8);
        std::memcpy(&pct[g], &p, 8);
        rate_f64[g] = sl->rate<double>(period);
        rate_i64[g] = zero_duration ? 0 : sl->rate<int64_t>(period);
        avg_f64[g] = sl->avg<double>();
        avg_i64[g] = 0;
      } else {
        auto sl = slice->series<int64_t>(static_cast<int>(s));
        if (!sl) return -EIO;
        diff[g] = static_cast<uint64_t>(sl->diff());
        pct[g] = static_cast<uint64_t>(sl->percentile(percentage));
        rate_f64[g] = sl->rate<double>(period);
        rate_i64[g] = zero_duration ? 0 : sl->rate<int64_t>(period);
        avg_f64[g] = sl->avg<double>();
        avg_i64[g] = sl->avg<int64_t>();
      }
    }
  }
  return 0;
}

// Back-pressure: the REAL CpuCountSampleGenerator with an output ring that has room for exactly
// `rows_budget` more rows (the ring is filled with dummy rows to the brim, then rows_budget of
// them are consumed) and no drop-if-full: the SAMPLE whose row does not fit makes
// handleRecordSample return -ENOSPC (PerCpuCountSampleGenerator.h:182-195) and consume() stop in
// front of it (CpuEventsGroup.h:1441-1444).  Outputs as hbtref_sample_deltas.
extern "C" int hbtref_sample_deltas_budget(const uint8_t* ring, uint64_t data_size, uint64_t tail, uint64_t head,
                                           uint32_t n_events, uint32_t max_records, uint64_t rows_budget,
                                           uint64_t* carry_tstamp, uint64_t* carry_last, uint64_t* rows,
                                           uint64_t rows_cap, hbt_cpu_result* res) {
  if (n_events == 0 || n_events > HBT_MAX_EVENTS) return -EINVAL;
  hbt_tsc_params tsc{0, 1, 0, 0};
  FakeMmapAlias mm;
  mm.init(ring, data_size, tail, head, tsc);
  auto confs = make_confs(n_events, 0);
  std::vector<std::string> names;
  for (auto& c : confs) names.push_back(c.id);
  hp::CpuCountSampleGenerator g(names, 0, -1, confs, rows_budget + 8, 0);
  attach(g, mm.page, tsc);
  load_counters(g, res->counters);
  g.count_.tstamp = *carry_tstamp;
  for (uint32_t i = 0; i < n_events; ++i) g.last_[i] = carry_last[i];
  g.num_unknown_count_interval_ = static_cast<uint32_t>(res->num_unknown_count_interval);
  const size_t row_bytes = g.kWriteByteSize;
  using Extra = hp::CpuCountSampleGenerator::RbExtraData;
  facebook::hbt::ringbuffer::Producer<Extra> filler(g.getOutput());
  facebook::hbt::ringbuffer::Consumer<Extra> cons(g.getOutput());
  std::vector<uint64_t> row(2 + n_events, 0xD0D0D0D0D0D0D0D0ull);
  uint64_t dummies = 0;
  while (filler.write(row.data(), row_bytes) >= 0) ++dummies;
  if (dummies < rows_budget) { detach(g); return -ERANGE; }
  for (uint64_t k = 0; k < rows_budget; ++k)
    if (cons.copy(row_bytes, row.data()) < 0) { detach(g); return -EIO; }
  dummies -= rows_budget;

  int64_t ret = run_consume(g, max_records);
  finish_result(g, ring, data_size, tail, mm.hdr, ret, res);
  uint64_t n = 0, seen = 0;
  int rc = 0;
  while (cons.copy(row_bytes, row.data()) >= 0) {
    if (seen++ < dummies) continue;
    if (n >= rows_cap) { rc = -ENOSPC; break; }
    std::memcpy(rows + n * (2 + n_events), row.data(), row_bytes);
    ++n;
  }
  res->n_rows = n;
  res->num_unknown_count_interval = g.num_unknown_count_interval_;
  res->carry_tstamp = g.count_.tstamp;
  *carry_tstamp = g.count_.tstamp;
  for (uint32_t i = 0; i < n_events; ++i) carry_last[i] = g.last_[i];
  detach(g);
  return rc;
}
