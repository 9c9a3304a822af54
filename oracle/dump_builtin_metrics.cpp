// TEST INFRASTRUCTURE -- golden-vector generator, not part of the product path.
//
// Prints the reference's own metric descriptors (makeAvailableMetrics,
// hbt/src/perf_event/BuiltinMetrics.cpp:470-1175) as JSON: for every metric id and every CPU
// architecture it has event references for, the events IN ORDER (= the index of their value in
// GroupReadValues::T::count[] / Sampling::Sample::value[], PerCpuGeneratorsTest.cpp:149-156) with
// nickname, PMU type, event id and the TscToKernelTime value transform (PmuEvent.h:218-226).  tests/golden/make_builtin_metrics.py builds this file
// against the reference sources where they lie (/root/reference) and commits the output as
// tests/golden/builtin_metrics.json, which pins the static table in
// dynolog-1_b200/csrc/metric_table.h.
#include <iostream>
#include <sstream>
#include <string>

#include "hbt/src/perf_event/BuiltinMetrics.h"
#include "hbt/src/perf_event/Metrics.h"

namespace hp = facebook::hbt::perf_event;

static std::string esc(const std::string& s) {
  std::string o;
  for (char c : s) {
    if (c == '"' || c == '\\') o += '\\';
    o += c;
  }
  return o;
}

int main() {
  auto metrics = hp::makeAvailableMetrics();
  std::cout << "{\n";
  bool first_m = true;
  for (const auto& [id, desc] : metrics->getMetricDescs()) {
    std::cout << (first_m ? "" : ",\n") << "  \"" << esc(id) << "\": {";
    first_m = false;
    bool first_a = true;
    for (const auto& [arch, refs] : desc->event_refs_by_arch) {
      std::ostringstream an;
      if (arch.has_value()) an << *arch; else an << "any";
      std::cout << (first_a ? "" : ", ") << "\"" << esc(an.str()) << "\": [";
      first_a = false;
      bool first_e = true;
      for (const auto& r : refs) {
        std::cout << (first_e ? "" : ", ") << "{\"nickname\": \"" << esc(r.nickname) << "\", \"pmu_type\": "
                  << static_cast<int>(r.pmu_type) << ", \"event_id\": \"" << esc(r.event_id)
                  << "\", \"tsc_to_kernel_time\": "
                  << ((r.transforms.flags & hp::EventValueTransforms::TscToKernelTime) ? 1 : 0) << "}";
        first_e = false;
      }
      std::cout << "]";
    }
    std::cout << "}";
  }
  std::cout << "\n}\n";
  return 0;
}
