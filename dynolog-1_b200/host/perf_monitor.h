// dynolog::Logger / JsonLogger / PerfMonitor over the GPU path: same interface and
// log keys as dynolog/src/Logger.h:24-45, dynolog/src/Logger.cpp:38-58 and
// dynolog/src/PerfMonitor.{h,cpp}.
#pragma once
#include <chrono>
#include <cstdio>
#include <map>
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

  // What dynolog logs per metric (PerfMonitor.cpp:38-73): the cumulative scaled count of every
  // event under its nickname, except the two metrics whose single event is reported as a rate
  // per millisecond of running time.  Those rates were computed on the GPU with the read
  // (GroupReadValues::getRatePerMs); nothing is divided here.
  struct RateKey {
    const char* metric_and_event;  // the metric id, whose only event carries the same nickname
    const char* log_key;
  };
  static constexpr RateKey kRateKeys[2] = {{"instructions", "mips"}, {"cycles", "mega_cycles_per_second"}};

  void log(Logger& logger) {
    for (const auto& [metric, reader] : countReaders_) {
      if (!reader) continue;
      const auto refs = reader->metric_desc->getEventRefs();
      const auto& values = readValues_[metric];
      if (!refs || !values) continue;  // the reference logs an error and moves on (:46-51)
      const char* rate_key = nullptr;
      for (const RateKey& k : kRateKeys)
        if (metric == k.metric_and_event) rate_key = k.log_key;
      for (size_t e = 0; e < refs->size(); ++e) {
        const std::string& nick = (*refs)[e].nickname;
        if (rate_key && nick == metric) logger.logFloat(rate_key, values->getRatePerMs(e));
        else logger.logUint(nick, values->getCount(e));
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

}  // namespace dynolog_gpu
