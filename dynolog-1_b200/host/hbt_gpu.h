// Host-side mirror of the reference's interface for the perf_event decode/aggregate
// path, written in the reference's language (C++17) on top of the C-ABI of
// include/hbt_b200.h.  Same class names, argument meaning and error behaviour as
//   hbt::perf_event::GroupReadValues      hbt/src/perf_event/CpuEventsGroup.h:368-535
//   hbt::perf_event::PerCpuCountReader    hbt/src/perf_event/PerCpuCountReader.h:56-149
//   hbt::perf_event::PerCpuBase::read     hbt/src/perf_event/PerCpuBase.h:104-119
//   hbt::perf_event::RbDataSlices         hbt/src/perf_event/CpuEventsGroup.h:44-63
//   hbt::mon::Monitor                     hbt/src/mon/Monitor.h:40-99,213-237,281-304,484-732
//   thread-switch generator               hbt/src/perf_event/PerCpuThreadSwitchGenerator.h:52-135,168-520
//   count-sample row sink + consumers     hbt/src/perf_event/PerCpuCountSampleGenerator.h:69,182-195,
//                                         hbt/src/ringbuffer/Producer.h:137-197, hbt/src/mon/MonData.h:69-128,
//                                         hbt/src/mon/TraceMonitor.h:34-109, TraceMonitor.cpp:25-100
// so that dynolog::PerfMonitor-shaped code (perf_monitor.h) compiles against it
// unchanged in shape.  All arithmetic of the path runs on the GPU: there is no
// host implementation behind these classes, and constructing a GpuContext
// without a usable device throws.
#pragma once
#include <algorithm>
#include <cerrno>
#include <cstdint>
#include <cstring>
#include <deque>
#include <map>
#include <memory>
#include <mutex>
#include <optional>
#include <set>
#include <stdexcept>
#include <string>
#include <system_error>
#include <vector>

#include "../../include/hbt_b200.h"

namespace hbt_gpu {

using CpuId = unsigned;
using TimeStamp = uint64_t;
constexpr unsigned kMaxCpus = HBT_MAX_CPUS;                 // hbt/src/common/System.h:195
constexpr TimeStamp kInvalidTimeStamp = ~TimeStamp(0);      // hbt/src/common/System.h:97

// hbt::CpuSet (hbt/src/common/System.h:198-283), reduced to what the path uses.
struct CpuSet {
  std::set<CpuId> cpus;
  CpuId max_cpu_id = 0;
  template <class TCont>
  static CpuSet makeFromCont(const TCont& c) {
    CpuSet s;
    for (auto cpu : c) {
      if (cpu < 0 || static_cast<unsigned>(cpu) >= kMaxCpus)
        throw std::invalid_argument("Invalid CPU ID: " + std::to_string(cpu));  // HBT_THROW_EINVAL
      s.cpus.insert(static_cast<CpuId>(cpu));
    }
    if (!s.cpus.empty()) s.max_cpu_id = *s.cpus.rbegin();
    return s;
  }
  static CpuSet makeRange(unsigned n) {
    std::vector<int> v(n);
    for (unsigned i = 0; i < n; ++i) v[i] = (int)i;
    return makeFromCont(v);
  }
  bool hasCpu(CpuId c) const { return cpus.count(c) != 0; }
  size_t numCpus() const { return cpus.size(); }
};

// RAII owner of one hbt_ctx.  Errors follow hbt's conventions: invalid arguments
// -> std::invalid_argument (HBT_ARG_CHECK*), anything the system refuses ->
// std::system_error (HBT_THROW_SYSTEM*), hbt/src/common/Defs.h:91-157.
class GpuContext {
 public:
  explicit GpuContext(int device = 0, uint32_t max_cpus = kMaxCpus, uint32_t max_events = HBT_MAX_EVENTS) {
    int rc = hbt_create(&ctx_, device, max_cpus, max_events);
    if (rc == -EINVAL) throw std::invalid_argument("hbt_create: invalid limits");
    if (rc < 0)
      throw std::system_error(-rc, std::generic_category(),
                              "hbt_create: no usable sm_100 CUDA device (there is no CPU fallback)");
  }
  ~GpuContext() { hbt_destroy(ctx_); }
  GpuContext(const GpuContext&) = delete;
  GpuContext& operator=(const GpuContext&) = delete;
  hbt_ctx* get() const { return ctx_; }
  void check(int64_t rc, const char* what) const {
    if (rc >= 0) return;
    const std::string msg = std::string(what) + ": " + hbt_last_error(ctx_);
    if (rc == -EINVAL) throw std::invalid_argument(msg);
    throw std::system_error((int)-rc, std::generic_category(), msg);
  }

 private:
  hbt_ctx* ctx_ = nullptr;
};

namespace perf_event {

struct DataSlice {
  const void* from;
  size_t size;
};
struct RbDataSlices {
  RbDataSlices(const void* p1, size_t n1) : first{p1, n1} {}
  RbDataSlices(const void* p1, size_t n1, const void* p2, size_t n2)
      : first{p1, n1}, second{DataSlice{p2, n2}} {}
  DataSlice first;
  std::optional<DataSlice> second;
};

// GroupReadValues<TMode>: the kernel's read_format POD plus the accessors of the
// reference.  getCount()'s multiplex scaling is what the GPU computed for this
// read (scaled_), not a host re-computation.
class GroupReadValues {
 public:
  explicit GroupReadValues(size_t n_events) : t_(3 + n_events, 0), scaled_(n_events, 0), rates_(n_events, 0.f) {
    if (n_events == 0) throw std::invalid_argument("n_events must be > 0");  // HBT_ARG_CHECK_GT
    t_[0] = n_events;
  }
  size_t getNumEvents() const { return (size_t)t_[0]; }
  size_t getNumReadBytes() const { return 8 * t_.size(); }
  uint64_t getTimeEnabled() const { return t_[1]; }
  uint64_t getTimeRunning() const { return t_[2]; }
  uint64_t getRawCount(size_t i) const { return t_.at(3 + checked(i)); }
  uint64_t getCount(size_t i) const { return scaled_.at(checked(i)); }
  std::vector<uint64_t> getRawCounts() const { return {t_.begin() + 3, t_.end()}; }
  std::vector<uint64_t> getCounts() const { return scaled_; }
  // float(double(getCount(i)) * 1e3 / double(getTimeRunning())): PerfMonitor.cpp:56-57
  float getRatePerMs(size_t i) const { return rates_.at(checked(i)); }
  bool hasMultiplexing() const { return t_[1] != 0 && t_[2] != t_[1]; }
  // Accessors on the summed triple, formulas as written in the reference
  // (CpuEventsGroup.h:481-501; note that getRunningRatio divides enabled by running)
  float getRunningRatio() const { return t_[1] == 0 ? 1.0f : (float)t_[1] / (float)t_[2]; }
  template <size_t scale = 100>
  uint64_t getRunningRatioInteger() const { return t_[1] == 0 ? scale : (scale * t_[2]) / t_[1]; }
  uint64_t* raw() { return t_.data(); }
  const uint64_t* raw() const { return t_.data(); }
  uint64_t* scaled() { return scaled_.data(); }
  float* rates() { return rates_.data(); }

 private:
  size_t checked(size_t i) const {
    if (i >= getNumEvents()) throw std::invalid_argument("Index out of range");  // HBT_ARG_CHECK_LT
    return i;
  }
  std::vector<uint64_t> t_;
  std::vector<uint64_t> scaled_;
  std::vector<float> rates_;
};

struct EventRef {
  std::string nickname;
};
// MetricDesc reduced to the event order / nicknames the path needs
// (hbt/src/perf_event/Metrics.h:122-179).
struct MetricDesc {
  std::string id;
  std::vector<EventRef> event_refs;
  std::optional<std::vector<EventRef>> getEventRefs() const {
    if (event_refs.empty()) return std::nullopt;
    return event_refs;
  }
  std::vector<std::string> eventNicknames() const {
    std::vector<std::string> v;
    for (auto& e : event_refs) v.push_back(e.nickname);
    return v;
  }
  std::optional<size_t> getNumEvents() const {
    if (event_refs.empty()) return std::nullopt;
    return event_refs.size();
  }
};

// Where the raw read_format bytes of each CPU's group leader come from: the
// ::read(2) of CpuEventsGroupBase::read (CpuEventsGroup.h:596-615) in a live
// daemon, a recorded trace in replay.  Returns false for a failed read.
class CounterSource {
 public:
  virtual ~CounterSource() = default;
  virtual bool readCpu(CpuId cpu, uint64_t* t, size_t n_events) = 0;
  virtual void nextInterval() {}
};

class TraceReplaySource : public CounterSource {
 public:
  // reads[interval][cpu_index][3+E], cpu_index = rank of the CPU id in the set
  TraceReplaySource(std::vector<uint64_t> reads, size_t n_cpus, size_t n_events)
      : reads_(std::move(reads)), n_cpus_(n_cpus), W_(3 + n_events) {}
  bool readCpu(CpuId cpu_index, uint64_t* t, size_t n_events) override {
    if (interval_ * n_cpus_ * W_ + (cpu_index + 1) * W_ > reads_.size()) return false;
    const uint64_t* src = &reads_[(interval_ * n_cpus_ + cpu_index) * W_];
    if (src[0] != n_events) return false;
    std::memcpy(t, src, 8 * W_);
    return true;
  }
  void nextInterval() override { ++interval_; }

 private:
  std::vector<uint64_t> reads_;
  size_t n_cpus_, W_, interval_ = 0;
};

// PerCpuCountReader: one counting-mode event group per CPU, read() = sum over CPUs
// (PerCpuBase::read) then scaling -- both on the GPU through the C-ABI.
class PerCpuCountReader {
 public:
  using ReadValues = GroupReadValues;

  PerCpuCountReader(const CpuSet& mon_cpus, std::shared_ptr<const MetricDesc> metric_desc_in,
                    std::shared_ptr<CounterSource> source, std::shared_ptr<GpuContext> gpu)
      : metric_desc{std::move(metric_desc_in)}, mon_cpus_{mon_cpus}, source_{std::move(source)}, gpu_{std::move(gpu)} {
    if (!metric_desc || !source_ || !gpu_) throw std::invalid_argument("null dependency");
    if (mon_cpus_.numCpus() == 0) throw std::invalid_argument("empty CpuSet");
    if (!metric_desc->getNumEvents()) throw std::invalid_argument("No events in metric_desc");
    staging_.assign(mon_cpus_.numCpus() * (3 + getNumEvents()), 0);
  }

  void open(bool /*pinned*/ = false) { open_ = true; }
  void close() {
    if (!open_) return;
    disable();
    open_ = false;
  }
  void enable() {
    if (!open_) throw std::invalid_argument("Cannot enable events that are not open");  // :1241
    enabled_ = true;
  }
  void disable() {
    if (!open_) throw std::invalid_argument("Cannot disable events that are not open");
    enabled_ = false;
  }
  bool isOpen() const { return open_; }
  bool isEnabled() const { return enabled_; }
  size_t getNumEvents() const { return *metric_desc->getNumEvents(); }
  ReadValues makeReadValues() const { return ReadValues{getNumEvents()}; }
  CpuSet getMonCpus() const { return mon_cpus_; }

  // PerCpuBase::read: false when not open or when any CPU's read fails.
  bool read(ReadValues& rv) const {
    if (!isOpen()) return false;
    if (rv.getNumEvents() != getNumEvents())
      throw std::invalid_argument("GroupReadValues of " + std::to_string(rv.getNumEvents()) + " but have " +
                                  std::to_string(getNumEvents()));  // CpuEventsGroup.h:599-601
    const size_t E = getNumEvents(), W = 3 + E, C = mon_cpus_.numCpus();
    size_t k = 0;
    for (CpuId cpu : mon_cpus_.cpus) {
      (void)cpu;
      uint64_t* t = &staging_[k * W];
      if (!source_->readCpu((CpuId)k, t, E)) t[0] = 0;  // marks the failed read for the kernel
      ++k;
    }
    source_->nextInterval();
    std::vector<uint64_t> sums(W), scaled(E);
    std::vector<float> rates(E);
    gpu_->check(hbt_counts_pipeline_host(gpu_->get(), staging_.data(), 1, (uint32_t)C, (uint32_t)E, nullptr, 0,
                                         sums.data(), scaled.data(), rates.data(), nullptr, nullptr),
                "hbt_counts_pipeline_host");
    if (sums[0] == 0) return false;  // some CPU's read failed -> std::nullopt upstream
    std::memcpy(rv.raw(), sums.data(), 8 * W);
    std::memcpy(rv.scaled(), scaled.data(), 8 * E);
    std::memcpy(rv.rates(), rates.data(), 4 * E);
    return true;
  }

  std::optional<ReadValues> read() const {
    auto rv = makeReadValues();
    if (read(rv)) return std::make_optional(rv);
    return std::nullopt;
  }

  const std::shared_ptr<const MetricDesc> metric_desc;  // public, PerfMonitor::log reads it

 private:
  CpuSet mon_cpus_;
  std::shared_ptr<CounterSource> source_;
  std::shared_ptr<GpuContext> gpu_;
  mutable std::vector<uint64_t> staging_;
  bool open_ = false, enabled_ = false;
};

// Batch replacement of the per-CPU CpuCountSampleGenerator::consume() loop
// (PerCpuSampleGeneratorBase::accumUntil, PerCpuSampleGeneratorBase.h:70-96): the
// slices every CPU's onCpuDataBufferRead hands out are submitted, then one
// consume() decodes all CPUs.
class PerCpuCountSampleDecoder {
 public:
  PerCpuCountSampleDecoder(const CpuSet& mon_cpus, size_t n_events, size_t data_size,
                           std::shared_ptr<GpuContext> gpu)
      : mon_cpus_{mon_cpus}, n_events_{n_events}, gpu_{std::move(gpu)} {
    if (n_events == 0 || n_events > HBT_MAX_EVENTS) throw std::invalid_argument("kNumEvents out of range");
    gpu_->check(hbt_ring_configure(gpu_->get(), mon_cpus_.max_cpu_id + 1, data_size), "hbt_ring_configure");
    results_.resize(mon_cpus_.max_cpu_id + 1);
  }
  // enable(): counters reset, count_.tstamp = now(), last_[] = 0 (enableImpl, :91-106)
  void enable(TimeStamp now) { gpu_->check(hbt_ring_reset_state(gpu_->get(), now), "hbt_ring_reset_state"); }
  // OnRbReadCallback body (CpuEventsGroup.h:63): tail = data_tail the slices start at
  void onCpuDataBufferRead(CpuId cpu, const RbDataSlices& s, uint64_t tail, uint64_t head) {
    if (!mon_cpus_.hasCpu(cpu)) throw std::invalid_argument("CPU " + std::to_string(cpu) + " not monitored");
    size_t n2 = 0;
    if (s.second) n2 = std::min(s.second->size, (size_t)(head - tail) - s.first.size);  // :1561 over-reports
    gpu_->check(hbt_submit_ring(gpu_->get(), cpu, s.first.from, s.first.size, s.second ? s.second->from : nullptr,
                                n2, tail),
                "hbt_submit_ring");
  }
  // Back-pressure: rows every CPU's output ring still takes.  consume() then stops a CPU with
  // -ENOSPC in front of the SAMPLE whose row would not fit, as the generator does when its ring is
  // full and nothing may be dropped (PerCpuCountSampleGenerator.h:182-195).
  void setRowsBudget(const std::map<CpuId, uint64_t>& free_rows) {
    std::vector<uint64_t> b(results_.size(), ~0ull);
    for (const auto& [cpu, n] : free_rows) b.at(cpu) = n;
    gpu_->check(hbt_set_rows_budget(gpu_->get(), b.data(), (uint32_t)b.size()), "hbt_set_rows_budget");
  }
  void clearRowsBudget() { gpu_->check(hbt_set_rows_budget(gpu_->get(), nullptr, 0), "hbt_set_rows_budget"); }
  // consume(max_num_records) for every CPU; returns what each CPU's consume() would.
  std::map<CpuId, int64_t> consume(unsigned max_num_records) {
    hbt_decode_desc d{};
    d.handled_mask = HBT_MODE_SAMPLING;
    d.sample_layout = HBT_SAMPLE_GROUP;
    d.aggregate = HBT_AGG_COUNT_DELTAS;
    d.n_events = (uint32_t)n_events_;
    d.max_records = max_num_records;
    d.tsc.time_mult = 1;
    gpu_->check(hbt_decode(gpu_->get(), &d), "hbt_decode");
    gpu_->check(hbt_fetch_results(gpu_->get(), results_.data(), (uint32_t)results_.size()), "hbt_fetch_results");
    rows_valid_ = false;
    std::map<CpuId, int64_t> ret;
    for (CpuId cpu : mon_cpus_.cpus) ret[cpu] = results_[cpu].ret;
    return ret;
  }
  // rows {tstamp, duration, deltas[n]} of the last consume(), kWriteByteSize each (:69).  The
  // first call after a consume() brings the rows of ALL CPUs back with one copy (the library
  // packs them on the device, hbt_fetch_rows_all); later calls slice that.
  std::vector<uint64_t> rows(CpuId cpu) {
    const auto& r = results_.at(cpu);
    const size_t w = 2 + n_events_;
    if (!rows_valid_) {
      uint64_t total = 0;
      for (const auto& x : results_) total += x.n_rows;
      rows_all_.resize(total * w + 1);
      rows_off_.assign(results_.size() + 1, 0);
      gpu_->check(hbt_fetch_rows_all(gpu_->get(), rows_all_.data(), total, rows_off_.data()), "hbt_fetch_rows_all");
      rows_valid_ = true;
    }
    const uint64_t* b = rows_all_.data() + rows_off_[cpu] * w;
    return std::vector<uint64_t>(b, b + r.n_rows * w);
  }
  uint64_t getNumRecordLost(CpuId c) const { return results_.at(c).counters[HBT_CNT_LOST]; }
  uint64_t getNumRecordSample(CpuId c) const { return results_.at(c).counters[HBT_CNT_SAMPLE]; }
  uint64_t getNumUnknownCountInterval(CpuId c) const { return results_.at(c).num_unknown_count_interval; }
  TimeStamp getLastTimeStamp(CpuId c) const { return results_.at(c).carry_tstamp; }
  uint64_t dataTail(CpuId c) const { return results_.at(c).new_tail; }

 private:
  CpuSet mon_cpus_;
  size_t n_events_;
  std::shared_ptr<GpuContext> gpu_;
  std::vector<hbt_cpu_result> results_;
  std::vector<uint64_t> rows_all_, rows_off_;
  bool rows_valid_ = false;
};

// tagstack::Event (hbt/src/tagstack/Event.h:28-227) is hbt_switch_event byte for byte.
using Event = hbt_switch_event;

// The virtual-id half of PerThreadStats (PerCpuThreadSwitchGenerator.h:52-135) as the
// generator's handlers drive it (:228-412).  The GPU tags thread events with the raw tid
// and reports COMM records as pseudo-events; apply() replays the bookkeeping over one
// CPU's events in stream order -- selectNextVid / createThreadInfo / registerExit, and
// active_tids.clear() on WriteErrorsStart -- rewrites the tags to vids and drops the
// pseudo-events.  Thread names, cmdlines and procfs reads stay with the caller.
class ThreadSwitchVidMapper {
 public:
  uint32_t next_vid = 1;
  std::map<uint32_t, uint32_t> active_tids;  // tid -> vid
  size_t apply(std::vector<Event>& ev) {
    size_t out = 0;
    for (size_t i = 0; i < ev.size(); ++i) {
      Event e = ev[i];
      if (e.type == HBT_SWEV_WRITE_ERRORS_START) active_tids.clear();  // :253-256
      const bool thread_ev = e.type == HBT_SWEV_COMM || e.type == HBT_SWEV_THREAD_CREATION ||
                             e.type == HBT_SWEV_THREAD_DESTRUCTION || e.type == HBT_SWEV_SWITCH_IN;
      if (thread_ev) {
        const uint32_t tid = e.tag;
        uint32_t vid = 0;
        if (tid != 0) {  // tid 0 is not remapped (:91-97)
          auto it = active_tids.find(tid);
          if (it == active_tids.end()) {
            vid = next_vid++;  // createThreadInfo / registerFork
            active_tids[tid] = vid;
          } else {
            vid = it->second;
          }
        }
        if (e.type == HBT_SWEV_THREAD_DESTRUCTION) active_tids.erase(tid);  // registerExit
        if (e.type == HBT_SWEV_COMM) continue;
        e.tag = vid;
      }
      ev[out++] = e;
    }
    ev.resize(out);
    return out;
  }
};

// Batch replacement of PerCpuThreadSwitchGenerator (PerCpuThreadSwitchGenerator.h:457-520):
// every CPU's context-switch ring is decoded in one call; events(cpu) is what that CPU's
// CpuThreadSwitchGenerator would have written to its output ring, vids included when the
// CPUs are read in ascending order (the order accumUntil visits them).
class PerCpuThreadSwitchDecoder {
 public:
  PerCpuThreadSwitchDecoder(const CpuSet& mon_cpus, uint8_t tid_level, size_t data_size,
                            std::shared_ptr<GpuContext> gpu)
      : mon_cpus_{mon_cpus}, tid_level_{tid_level}, gpu_{std::move(gpu)} {
    gpu_->check(hbt_ring_configure(gpu_->get(), mon_cpus_.max_cpu_id + 1, data_size), "hbt_ring_configure");
    results_.resize(mon_cpus_.max_cpu_id + 1);
  }
  void onCpuDataBufferRead(CpuId cpu, const RbDataSlices& s, uint64_t tail, uint64_t head) {
    if (!mon_cpus_.hasCpu(cpu)) throw std::invalid_argument("CPU " + std::to_string(cpu) + " not monitored");
    size_t n2 = 0;
    if (s.second) n2 = std::min(s.second->size, (size_t)(head - tail) - s.first.size);
    gpu_->check(hbt_submit_ring(gpu_->get(), cpu, s.first.from, s.first.size, s.second ? s.second->from : nullptr,
                                n2, tail),
                "hbt_submit_ring");
  }
  std::map<CpuId, int64_t> consume(unsigned max_num_records) {
    hbt_decode_desc d{};
    d.handled_mask = HBT_MODE_CONTEXT_SWITCH;
    d.aggregate = HBT_AGG_SWITCH_EVENTS;
    d.max_records = max_num_records;
    d.tsc.time_mult = 1;
    d.tid_level = tid_level_;
    gpu_->check(hbt_decode(gpu_->get(), &d), "hbt_decode");
    gpu_->check(hbt_fetch_results(gpu_->get(), results_.data(), (uint32_t)results_.size()), "hbt_fetch_results");
    std::map<CpuId, int64_t> ret;
    events_.clear();
    for (CpuId cpu : mon_cpus_.cpus) {  // ascending: the vid map is shared by all CPUs
      ret[cpu] = results_[cpu].ret;
      std::vector<Event> ev(results_[cpu].n_rows);
      if (!ev.empty())
        gpu_->check(hbt_fetch_switch_events(gpu_->get(), cpu, ev.data(), ev.size()), "hbt_fetch_switch_events");
      thread_stats_.apply(ev);
      events_[cpu] = std::move(ev);
    }
    return ret;
  }
  const std::vector<Event>& events(CpuId cpu) const { return events_.at(cpu); }
  size_t getNumEntries(CpuId cpu) const { return events_.at(cpu).size(); }
  TimeStamp getLastTimeStamp(CpuId c) const { return results_.at(c).carry_tstamp; }
  uint64_t getNumRecordLost(CpuId c) const { return results_.at(c).counters[HBT_CNT_LOST]; }
  uint64_t getNumRecordSwitchCpuWide(CpuId c) const { return results_.at(c).counters[HBT_CNT_SWITCH_CPU_WIDE]; }
  uint64_t dataTail(CpuId c) const { return results_.at(c).new_tail; }
  const ThreadSwitchVidMapper& threadStats() const { return thread_stats_; }

 private:
  CpuSet mon_cpus_;
  uint8_t tid_level_;
  std::shared_ptr<GpuContext> gpu_;
  std::vector<hbt_cpu_result> results_;
  std::map<CpuId, std::vector<Event>> events_;
  ThreadSwitchVidMapper thread_stats_;
};

// What a CpuCountSampleGenerator's output ring is to its consumers: a byte FIFO of
// nextPow2(rb_min_num_entries * kWriteByteSize) bytes (hbt/src/ringbuffer/RingBuffer.h:69)
// holding rows {tstamp, duration, deltas[n]}.  write() follows handleRecordSample's
// tail (PerCpuCountSampleGenerator.h:182-195): a row that does not fit is refused with
// -ENOSPC unless num_counts_drop_if_full > 0, in which case that many of the OLDEST rows
// are dropped first (Producer::dropN, Producer.h:181-197) and the write is retried.
// The GPU decode produces a whole batch of rows at once, so unlike the reference a
// refused row cannot stop consume() before its record: size the sink for a batch or
// set the drop rule -- write() reports how many rows it had to refuse.
class CountSampleSink {
 public:
  CountSampleSink(size_t n_events, size_t rb_min_num_entries, size_t num_counts_drop_if_full)
      : kNumEvents{n_events}, kWriteByteSize{sizeof(uint64_t) * (2 + n_events)},
        kNumBytesDropIfFull{num_counts_drop_if_full * kWriteByteSize},
        kDataPoolByteSize{nextPow2(rb_min_num_entries * kWriteByteSize)} {
    if (kNumBytesDropIfFull > kDataPoolByteSize) throw std::invalid_argument("drop size larger than the buffer");
  }
  const size_t kNumEvents, kWriteByteSize, kNumBytesDropIfFull, kDataPoolByteSize;

  // One row; 0 or -ENOSPC.
  int write(const uint64_t* row) {
    if (usedBytes() + kWriteByteSize > kDataPoolByteSize) {
      if (kNumBytesDropIfFull == 0) return -ENOSPC;
      dropN(kNumBytesDropIfFull);
      if (usedBytes() + kWriteByteSize > kDataPoolByteSize) return -ENOSPC;
    }
    data_.insert(data_.end(), row, row + 2 + kNumEvents);
    return 0;
  }
  // n_rows rows in stream order; returns the number refused (0 when all were taken).
  size_t write(const uint64_t* rows, size_t n_rows) {
    size_t refused = 0;
    for (size_t r = 0; r < n_rows; ++r)
      if (write(rows + r * (2 + kNumEvents)) < 0) ++refused;
    return refused;
  }
  // Header::dropN: forget the oldest bytes (all of them if fewer are stored)
  void dropN(size_t n_bytes) {
    head_ += std::min(n_bytes / sizeof(uint64_t), data_.size() - head_);
    if (head_ * 2 > data_.size() && head_ >= 1024) {  // reclaim the consumed front now and then
      data_.erase(data_.begin(), data_.begin() + (std::ptrdiff_t)head_);
      head_ = 0;
    }
  }
  size_t getNumEntries() const { return (data_.size() - head_) / (2 + kNumEvents); }
  // rows write() still takes before it answers -ENOSPC (no drop-if-full)
  size_t freeRows() const { return (kDataPoolByteSize - usedBytes()) / kWriteByteSize; }
  size_t usedBytes() const { return (data_.size() - head_) * sizeof(uint64_t); }
  // Consumer side (ringbuffer::Consumer::readInTx / relSeekInTx / commitTx reduced to a cursor);
  // the returned row is contiguous and stays valid until the next write/consume.
  const uint64_t* peek(size_t i) const {
    return i < getNumEntries() ? data_.data() + head_ + i * (2 + kNumEvents) : nullptr;
  }
  void consume(size_t n_rows) { dropN(n_rows * kWriteByteSize); }

 private:
  static size_t nextPow2(size_t v) {
    size_t p = 1;
    while (p < v) p <<= 1;
    return p;
  }
  std::vector<uint64_t> data_;
  size_t head_ = 0;  // first live word
};

}  // namespace perf_event

namespace mon {

// hbt::mon::CountData (hbt/src/mon/MonData.h:69-128): rows of one compute unit.
struct CountData {
  const std::vector<std::string> event_names;
  const size_t kRowStride;  // in elements
  explicit CountData(const std::vector<std::string>& names) : event_names{names}, kRowStride{names.size() + 2} {}
  size_t numEvents() const noexcept { return event_names.size(); }
  void write(const uint64_t* d) { data.insert(data.end(), d, d + (2 + numEvents())); }
  size_t numIntervals() const { return data.size() / kRowStride; }
  // {first row's start, last row's start + duration}
  std::optional<std::pair<TimeStamp, TimeStamp>> getTimeStampRange() const {
    if (data.size() % kRowStride != 0) throw std::runtime_error("CountData: partial row");  // HBT_THROW_ASSERT_IF
    if (data.empty()) return std::nullopt;
    const size_t last = data.size() - (2 + numEvents());
    return std::make_pair(data[0], data[last] + data[last + 1]);
  }
  std::unique_ptr<CountData> filter(const std::vector<bool>& sel) const {
    auto cd = std::make_unique<CountData>(event_names);
    for (size_t r = 0, i = 0; r < data.size(); r += kRowStride, ++i)
      if (sel[i]) {
        std::vector<uint64_t> row(data.begin() + (std::ptrdiff_t)r, data.begin() + (std::ptrdiff_t)(r + kRowStride));
        cd->write(row.data());
      }
    return cd;
  }
  std::optional<size_t> findCountAt(TimeStamp tstamp) const {
    for (size_t r = 0, i = 0; r < data.size(); r += kRowStride, ++i)
      if (data[r] <= tstamp && data[r] + data[r + 1] > tstamp) return i;
    return std::nullopt;
  }
  std::deque<uint64_t> data;
};

// The count-generator half of hbt::mon::TraceMonitor (hbt/src/mon/TraceMonitor.h:34-109,
// TraceMonitor.cpp:25-100): per-CPU sinks fed by the GPU decoder's rows, consumed either
// through a callback or into CountData up to a stop timestamp.
class TraceMonitor {
 public:
  TraceMonitor(const CpuSet& mon_cpus, std::vector<std::string> event_nicknames, size_t rb_min_num_entries,
               size_t num_counts_drop_if_full)
      : mon_cpus_{mon_cpus}, nicknames_{std::move(event_nicknames)} {
    for (CpuId cpu : mon_cpus_.cpus)
      sinks_.emplace(cpu, perf_event::CountSampleSink(nicknames_.size(), rb_min_num_entries, num_counts_drop_if_full));
  }
  perf_event::CountSampleSink& sink(CpuId cpu) {
    auto it = sinks_.find(cpu);
    if (it == sinks_.end()) throw std::invalid_argument("CPU " + std::to_string(cpu) + " not monitored");
    return it->second;
  }
  // Tell the decoder how many rows every sink still takes, so that its next consume() stops where
  // the reference's would (sinks that drop their oldest rows when full never push back).
  void applyBackPressure(perf_event::PerCpuCountSampleDecoder& dec) {
    std::lock_guard<std::mutex> lock{counts_mutex_};
    std::map<CpuId, uint64_t> free_rows;
    for (CpuId cpu : mon_cpus_.cpus)
      if (sink(cpu).kNumBytesDropIfFull == 0) free_rows[cpu] = sink(cpu).freeRows();
    dec.setRowsBudget(free_rows);
  }
  // Move the rows of the decoder's last consume() into the sinks (the producer side:
  // what handleRecordSample's output_producer_.write does row by row).  Returns rows refused
  // (none after applyBackPressure).
  size_t accum(perf_event::PerCpuCountSampleDecoder& dec) {
    std::lock_guard<std::mutex> lock{counts_mutex_};
    size_t refused = 0;
    for (CpuId cpu : mon_cpus_.cpus) {
      const std::vector<uint64_t> rows = dec.rows(cpu);
      refused += sink(cpu).write(rows.data(), rows.size() / (2 + nicknames_.size()));
    }
    return refused;
  }
  // func(cpu, count_idx, event_nicknames, tstamp, duration, const uint64_t* counts, kNumEvents)
  // for every stored row, CPUs in ascending order, then the rows are consumed (TraceMonitor.h:45-109).
  template <class TFunc>
  void applyToCountSamplesAndConsume(TFunc& func) {
    std::lock_guard<std::mutex> lock{counts_mutex_};
    for (CpuId cpu : mon_cpus_.cpus) {
      auto& s = sink(cpu);
      const size_t n = s.getNumEntries();
      for (size_t i = 0; i < n; ++i) {
        const uint64_t* p = s.peek(i);
        func(cpu, (int)i, nicknames_, (TimeStamp)p[0], (TimeStamp)p[1], p + 2, nicknames_.size());
      }
      s.consume(n);
    }
  }
  // releaseCountData_ (TraceMonitor.cpp:25-100) without slices: rows whose interval ends
  // before stop_ts move into a CountData; the first row reaching stop_ts stays in the sink.
  std::unique_ptr<CountData> releaseCountData(CpuId cpu, TimeStamp stop_ts) {
    std::lock_guard<std::mutex> lock{counts_mutex_};
    auto& s = sink(cpu);
    auto cd = std::make_unique<CountData>(nicknames_);
    size_t n = 0;
    for (; n < s.getNumEntries(); ++n) {
      const uint64_t* p = s.peek(n);
      if (p[0] + p[1] >= stop_ts) break;
      cd->write(p);
    }
    s.consume(n);
    return cd;
  }

 private:
  CpuSet mon_cpus_;
  std::vector<std::string> nicknames_;
  std::map<CpuId, perf_event::CountSampleSink> sinks_;
  std::mutex counts_mutex_;
};

// hbt::mon::Monitor: element map, mux queue, Closed/Open/Enabled state machine,
// one mutex (hbt/src/mon/Monitor.h:40-99, 484-520, 681-732).
template <class MuxGroupIdType = std::optional<std::string>, class ElemIdType = std::string>
class Monitor {
  enum class State { Closed, Open, Enabled };

 public:
  using TCountReader = perf_event::PerCpuCountReader;
  using ElemId = ElemIdType;
  using MuxGroupId = MuxGroupIdType;

  bool open() { return lockedTo(State::Open); }
  bool close() { return lockedTo(State::Closed); }
  bool enable() { return lockedTo(State::Enabled); }
  bool disable() { return lockedTo(State::Open); }
  bool isOpen() const {
    std::lock_guard<std::mutex> l{mutex_};
    return state_ == State::Open;
  }
  bool isEnabled() const {
    std::lock_guard<std::mutex> l{mutex_};
    return state_ == State::Enabled;
  }
  void muxRotate() {
    std::lock_guard<std::mutex> l{mutex_};
    if (mux_queue_.empty()) return;
    std::rotate(mux_queue_.rbegin(), mux_queue_.rbegin() + 1, mux_queue_.rend());
    sync_();
  }

  std::shared_ptr<TCountReader> emplaceCountReader(const MuxGroupId& mux_group_id, const ElemId& elem_id,
                                                   std::shared_ptr<const perf_event::MetricDesc> metric_desc,
                                                   std::shared_ptr<perf_event::CounterSource> source,
                                                   const CpuSet& mon_cpus, std::shared_ptr<GpuContext> gpu) {
    std::lock_guard<std::mutex> l{mutex_};
    if (count_readers_.count(elem_id) != 0)  // HBT_ARG_CHECK_EQ, Monitor.h:291-293
      throw std::invalid_argument("There is already an interval source with key: \"" + elem_id + "\"");
    auto& g = mux_groups_[mux_group_id];
    if (g.empty()) mux_queue_.push_back(mux_group_id);
    g.insert(elem_id);
    auto [it, ok] = count_readers_.emplace(
        elem_id, std::make_shared<TCountReader>(mon_cpus, std::move(metric_desc), std::move(source), std::move(gpu)));
    (void)ok;
    sync_();
    return it->second;
  }

  std::map<ElemId, std::optional<typename TCountReader::ReadValues>> readAllCounts() const {
    std::lock_guard<std::mutex> l{mutex_};
    std::map<ElemId, std::optional<typename TCountReader::ReadValues>> rvs;
    for (auto& [k, cr] : count_readers_) rvs.emplace(k, cr->read());
    return rvs;
  }
  std::optional<typename TCountReader::ReadValues> readCounts(const ElemId& elem_id) const {
    std::lock_guard<std::mutex> l{mutex_};
    auto it = count_readers_.find(elem_id);
    if (it == count_readers_.end()) return std::nullopt;
    return it->second->read();
  }

 private:
  bool lockedTo(State s) {
    std::lock_guard<std::mutex> l{mutex_};
    if (state_ == s) return false;
    state_ = s;
    sync_();
    return true;
  }
  void sync_() {
    if (mux_queue_.empty()) return;
    for (auto& [k, p] : count_readers_) {
      auto& e = *p;
      switch (state_) {
        case State::Closed:
          if (e.isEnabled()) e.disable();
          if (e.isOpen()) e.close();
          break;
        case State::Open:
          if (e.isEnabled()) e.disable();
          if (!e.isOpen()) e.open();
          break;
        case State::Enabled:
          if (!e.isOpen()) e.open();
          // only the group at the front of the multiplexing queue is enabled
          if (mux_groups_.at(mux_queue_.front()).count(k) > 0) {
            if (!e.isEnabled()) e.enable();
          } else if (e.isEnabled()) {
            e.disable();
          }
          break;
      }
    }
  }
  mutable std::mutex mutex_;
  State state_ = State::Closed;
  std::map<ElemId, std::shared_ptr<TCountReader>> count_readers_;
  std::map<MuxGroupId, std::set<ElemId>> mux_groups_;
  std::deque<MuxGroupId> mux_queue_;
};

}  // namespace mon
}  // namespace hbt_gpu
