// dynolog::Logger / JsonLogger / PerfMonitor over the GPU path: same interface and
// log keys as dynolog/src/Logger.h:24-45, dynolog/src/Logger.cpp:38-58 and
// dynolog/src/PerfMonitor.{h,cpp}.
#pragma once
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <map>
#include <numeric>
#include <type_traits>
#include <ostream>
#include <string>

#include "hbt_gpu.h"

namespace dynolog_gpu {

class Logger {
 public:
  using Timestamp = std::chrono::time_point<std::chrono::system_clock>;
  virtual ~Logger() {}
  virtual void setTimestamp(Timestamp ts = std::chrono::system_clock::now()) = 0;
  virtual void logInt(const std::string& key, int64_t val) = 0;
  virtual void logFloat(const std::string& key, float val) = 0;
  virtual void logUint(const std::string& key, uint64_t val) = 0;
  virtual void logStr(const std::string& key, const std::string& val) = 0;
  virtual void finalize() = 0;
};

// JsonLogger without the nlohmann/glog dependencies: floats are formatted
// "{:.3f}" and stored as strings exactly like Logger.cpp:42-44; finalize() dumps
// one JSON object (keys sorted, like nlohmann's default map) and clears.
class JsonLogger : public Logger {
 public:
  explicit JsonLogger(std::ostream& os) : os_(os) {}
  void setTimestamp(Timestamp ts) override { ts_ = ts; }
  void logInt(const std::string& key, int64_t val) override { json_[key] = std::to_string(val); }
  void logFloat(const std::string& key, float val) override {
    char buf[64];
    std::snprintf(buf, sizeof(buf), "%.3f", val);
    json_[key] = std::string("\"") + buf + "\"";
  }
  void logUint(const std::string& key, uint64_t val) override { json_[key] = std::to_string(val); }
  void logStr(const std::string& key, const std::string& val) override { json_[key] = "\"" + val + "\""; }
  void finalize() override {
    os_ << "{";
    bool first = true;
    for (auto& [k, v] : json_) {
      os_ << (first ? "" : ",") << "\"" << k << "\":" << v;
      first = false;
    }
    os_ << "}\n";
    json_.clear();
  }
  const std::map<std::string, std::string>& sample() const { return json_; }

 private:
  std::ostream& os_;
  Timestamp ts_;
  std::map<std::string, std::string> json_;
};

using ElemId = hbt_gpu::mon::Monitor<>::ElemId;
using MuxGroupId = hbt_gpu::mon::Monitor<>::MuxGroupId;
using TCountReader = hbt_gpu::mon::Monitor<>::TCountReader;

// Registry of metric descriptors (the part of hbt::perf_event::Metrics the path uses).
struct Metrics {
  std::map<std::string, std::shared_ptr<const hbt_gpu::perf_event::MetricDesc>> descs;
  std::shared_ptr<const hbt_gpu::perf_event::MetricDesc> getMetricDesc(const std::string& id) const {
    auto it = descs.find(id);
    return it == descs.end() ? nullptr : it->second;
  }
};

class PerfMonitor {
 public:
  // sources: where each metric's per-CPU reads come from (perf fds live, a trace in replay)
  PerfMonitor(const hbt_gpu::CpuSet& monCpus, const std::vector<ElemId>& metricIds,
              std::shared_ptr<Metrics> availableMetrics,
              std::map<ElemId, std::shared_ptr<hbt_gpu::perf_event::CounterSource>> sources,
              std::shared_ptr<hbt_gpu::GpuContext> gpu)
      : monCpus_(monCpus), defaultMuxGroupId_() {
    for (const ElemId& id : metricIds) {
      countReaders_.emplace(id, hbtMon_.emplaceCountReader(defaultMuxGroupId_, id, availableMetrics->getMetricDesc(id),
                                                           sources.at(id), monCpus_, gpu));
    }
    hbtMon_.open();
    hbtMon_.enable();
  }

  void step() { readValues_ = hbtMon_.readAllCounts(); }

  // PerfMonitor::log (PerfMonitor.cpp:38-73): cumulative values; `instructions` and
  // `cycles` become mips / mega_cycles_per_second, everything else logUint(nickname).
  void log(Logger& logger) {
    for (auto& [id, countReader] : countReaders_) {
      if (countReader == nullptr) continue;
      auto eventRefsOpt = countReader->metric_desc->getEventRefs();
      auto readValueOpt = readValues_[id];
      if (!eventRefsOpt.has_value() || !readValueOpt.has_value()) continue;  // LOG(ERROR) + continue
      for (size_t i = 0; i < eventRefsOpt->size(); i++) {
        const std::string nickname = (*eventRefsOpt)[i].nickname;
        const uint64_t count = readValueOpt->getCount(i);
        if (id == "instructions" && nickname == "instructions") {
          logger.logFloat("mips", readValueOpt->getRatePerMs(i));
        } else if (id == "cycles" && nickname == "cycles") {
          logger.logFloat("mega_cycles_per_second", readValueOpt->getRatePerMs(i));
        } else {
          logger.logUint(nickname, count);
        }
      }
    }
  }

  const std::map<ElemId, std::optional<TCountReader::ReadValues>>& readValues() const { return readValues_; }

 protected:
  hbt_gpu::mon::Monitor<> hbtMon_;
  const hbt_gpu::CpuSet monCpus_;
  const MuxGroupId defaultMuxGroupId_;
  std::map<ElemId, std::optional<TCountReader::ReadValues>> readValues_;
  std::map<ElemId, std::shared_ptr<TCountReader>> countReaders_;
};

// dynolog::MetricSeries<T> (dynolog/src/metric_frame/MetricSeries.h:18-260): a ring of the
// last bufferLen samples with diff / rate / avg / percentile over [begin, end) given as
// distances from the oldest sample (the reference's Iterator is exactly such a distance).
// This is where the per-interval scaled counts (GroupReadValues::getCount) and the deltas
// k_scaled_diff produces end up when dynolog keeps metric tables.
template <typename T>
class MetricSeries final {
 public:
  MetricSeries(size_t bufferLen, std::string name, std::string description)
      : data_(bufferLen), name_{std::move(name)}, description_{std::move(description)} {}
  const std::string& name() const { return name_; }
  const std::string& description() const { return description_; }
  void addSample(const T& sample) {  // :148-158: overwrite the oldest when full
    data_[tail_] = sample;
    tail_ = (tail_ + 1) % data_.size();
    if (size_ == data_.size()) head_ = (head_ + 1) % data_.size();
    else ++size_;
  }
  size_t size() const { return size_; }
  size_t maxLength() const { return data_.size(); }
  const T& at(size_t idx) const { return data_[(head_ + idx) % data_.size()]; }
  const T& operator[](size_t idx) const { return at(idx); }
  size_t begin() const { return 0; }
  size_t end() const { return size_; }

  T diff(std::optional<size_t> b = std::nullopt, std::optional<size_t> e = std::nullopt) const {
    return at(e.value_or(end()) - 1) - at(b.value_or(begin()));  // :221-227
  }
  template <typename R>
  R rate(std::chrono::steady_clock::duration period, std::chrono::steady_clock::duration duration,
         std::optional<size_t> b = std::nullopt, std::optional<size_t> e = std::nullopt) const {
    const T value = diff(b, e);
    if constexpr (std::is_integral_v<R>) {  // :235-240
      if (period > duration) return value * (period / duration);
      return value / (duration / period);
    } else {                                  // :241-249
      return static_cast<R>(value) *
             static_cast<double>(std::chrono::duration_cast<std::chrono::microseconds>(period).count()) /
             std::chrono::duration_cast<std::chrono::microseconds>(duration).count();
    }
  }
  template <typename R>
  R avg(std::optional<size_t> b = std::nullopt, std::optional<size_t> e = std::nullopt) const {
    const size_t b_ = b.value_or(begin()), e_ = e.value_or(end());
    R acc{0};
    for (size_t i = b_; i < e_; ++i) acc += at(i);  // std::accumulate(begin_, end_, R{0})
    return acc / static_cast<int>(e_ - b_);        // Iterator::difference_type is int (:22)
  }
  T percentile(float percentage, std::optional<size_t> b = std::nullopt,
               std::optional<size_t> e = std::nullopt) const {
    const size_t b_ = b.value_or(begin()), e_ = e.value_or(end());
    std::vector<T> copy;
    for (size_t i = b_; i < e_; ++i) copy.push_back(at(i));
    const long nth = lround(percentage * (copy.size() - 1));  // :214
    std::nth_element(copy.begin(), copy.begin() + nth, copy.end());
    return copy[nth];
  }

 private:
  std::vector<T> data_;
  size_t head_ = 0, tail_ = 0, size_ = 0;
  const std::string name_, description_;
};

}  // namespace dynolog_gpu
