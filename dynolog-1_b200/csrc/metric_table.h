// Static table of metric descriptors: for every metric id -- and CPU architecture,
// where the reference registers per-architecture event lists -- the event nicknames IN ORDER
// (position = index of the event's value in GroupReadValues::T::count[] and Sampling::Sample::value[],
// hbt/src/perf_event/Metrics.h:122-179, PerCpuGeneratorsTest.cpp:149-156), as
// makeAvailableMetrics registers them (hbt/src/perf_event/BuiltinMetrics.cpp:470-1175).
// Pinned against the reference's own registry: tests/golden/builtin_metrics.json is the output of
// oracle/dump_builtin_metrics.cpp compiled against the reference's BuiltinMetrics.cpp, and
// tests/test_metric_table.py compares this table with it entry by entry.
//
// The reference only DESCRIBES which events make up a metric; no code divides them.  The named
// ratios below are therefore build-defined (parity unpinned), in the one convention the reference
// has for a derived value, double(getCount(a)) / double(getCount(b)) (dynolog/src/PerfMonitor.cpp:56-57).
// The last two descriptors are build-defined as well: miss rates over the kernel's generic
// hardware-cache events, named the way the reference names them (BuiltinMetrics.cpp:26-54,299-316;
// aliases dTLB-load-misses / LLC-load-misses :460-464).
//
// Included by hbt_b200.cu only.  The ratio index pairs are also laid down in __constant__ memory
// (c_metric_ratios) so that a pipeline call selected by metric id uploads nothing.
#pragma once
#include "../../include/hbt_b200.h"

namespace hbt {

static const char* const kNick0[] = {"sys_cpu_clock", "user_cpu_clock"};
static const char* const kNick1[] = {"context_switches", "cycles", "instructions"};
static const char* const kRatioName1[] = {"ipc", "cycles_per_context_switch"};
static const hbt_ratio kRatio1[] = {{2, 1}, {1, 0}};
static const char* const kNick2[] = {"cycles"};
static const char* const kNick3[] = {"ref_cycles", "ucycles", "kcycles"};
static const char* const kRatioName3[] = {"user_cycle_fraction", "kernel_cycle_fraction"};
static const hbt_ratio kRatio3[] = {{1, 0}, {2, 0}};
static const char* const kNick4[] = {"cpu_clock", "instructions", "cycles", "context_switches", "page_faults_minor", "page_faults_major", "alignment_faults", "blocked_ns", "iowait_ns", "runtime_ns", "sleep_ns", "wait_ns"};
static const char* const kRatioName4[] = {"ipc"};
static const hbt_ratio kRatio4[] = {{1, 2}};
static const char* const kNick5[] = {"dram_access_reads"};
static const char* const kNick6[] = {"dram_access_reads"};
static const char* const kNick7[] = {"dram_access_reads"};
static const char* const kNick8[] = {"dram_access_reads"};
static const char* const kNick9[] = {"dram_access_reads"};
static const char* const kNick10[] = {"instr_dp_scalar", "instr_dp_128b_packed", "instr_dp_256b_packed", "instr_dp_512b_packed"};
static const char* const kNick11[] = {"instr_sp_scalar", "instr_sp_128b_packed", "instr_sp_256b_packed", "instr_sp_512b_packed"};
static const char* const kNick12[] = {"context_switches", "cpu_migrations", "page_faults"};
static const char* const kNick13[] = {"instructions"};
static const char* const kNick14[] = {"cycles", "instructions"};
static const char* const kRatioName14[] = {"ipc"};
static const hbt_ratio kRatio14[] = {{1, 0}};
static const char* const kNick15[] = {"l3_cache_misses", "instructions"};
static const char* const kRatioName15[] = {"l3_cache_misses_per_instruction"};
static const hbt_ratio kRatio15[] = {{0, 1}};
static const char* const kNick16[] = {"page_faults", "page_faults_minor", "page_faults_major"};
static const char* const kNick17[] = {"sys_enter_clone", "sys_enter_clone"};
static const char* const kNick18[] = {"unhalted", "retired", "unhalted_time", "issued", "recovery", "not_delivered"};
static const char* const kNick19[] = {"unhalted", "retired", "unhalted_time", "issued", "recovery", "not_delivered"};
static const char* const kNick20[] = {"mem_stalls", "l1_miss", "cycles"};
static const char* const kRatioName20[] = {"l1_bound_fraction"};
static const hbt_ratio kRatio20[] = {{0, 2}};
static const char* const kNick21[] = {"cycles", "l1_miss", "l2_miss"};
static const char* const kNick22[] = {"ifdata", "cycles"};
static const char* const kRatioName22[] = {"icache_stall_fraction"};
static const hbt_ratio kRatio22[] = {{0, 1}};
static const char* const kNick23[] = {"unhalted", "outstanding_bw", "outstanding_lat", "cycles"};
static const char* const kNick24[] = {"llc_load_misses", "llc_load_hits"};
static const char* const kRatioName24[] = {"llc_load_miss_rate"};
static const hbt_ratio kRatio24[] = {{0, 1}};
static const char* const kNick25[] = {"dtlb_load_misses", "dtlb_load_hits"};
static const char* const kRatioName25[] = {"dtlb_load_miss_rate"};
static const hbt_ratio kRatio25[] = {{0, 1}};

constexpr uint32_t kNumMetrics = 26;
constexpr uint32_t kMaxMetricRatios = 2;

static const hbt_metric_desc kMetricTable[kNumMetrics] = {
    {"cpu_clock", "any", 2, kNick0, 0, nullptr, nullptr, 1, 0, 0x0ull},
    {"cs_ipc", "any", 3, kNick1, 2, kRatio1, kRatioName1, 1, 1, 0x0ull},
    {"cycles", "any", 1, kNick2, 0, nullptr, nullptr, 1, 2, 0x0ull},
    {"cycles_breakdown", "any", 3, kNick3, 2, kRatio3, kRatioName3, 1, 3, 0x0ull},
    {"dqos", "any", 12, kNick4, 1, kRatio4, kRatioName4, 1, 4, 0x0ull},
    {"dram_access_reads", "BDW", 1, kNick5, 0, nullptr, nullptr, 1, 5, 0x0ull},
    {"dram_access_reads", "BDW-DE", 1, kNick6, 0, nullptr, nullptr, 1, 6, 0x0ull},
    {"dram_access_reads", "BDX", 1, kNick7, 0, nullptr, nullptr, 1, 7, 0x0ull},
    {"dram_access_reads", "SKL", 1, kNick8, 0, nullptr, nullptr, 1, 8, 0x0ull},
    {"dram_access_reads", "SKX", 1, kNick9, 0, nullptr, nullptr, 1, 9, 0x0ull},
    {"fp_instrs_double_precision", "any", 4, kNick10, 0, nullptr, nullptr, 1, 10, 0x0ull},
    {"fp_instrs_single_precision", "any", 4, kNick11, 0, nullptr, nullptr, 1, 11, 0x0ull},
    {"generic_sw", "any", 3, kNick12, 0, nullptr, nullptr, 1, 12, 0x0ull},
    {"instructions", "any", 1, kNick13, 0, nullptr, nullptr, 1, 13, 0x0ull},
    {"ipc", "any", 2, kNick14, 1, kRatio14, kRatioName14, 1, 14, 0x0ull},
    {"l3_cache_misses_per_instruction", "any", 2, kNick15, 1, kRatio15, kRatioName15, 1, 15, 0x0ull},
    {"page_faults", "any", 3, kNick16, 0, nullptr, nullptr, 1, 16, 0x0ull},
    {"system_calls", "any", 2, kNick17, 0, nullptr, nullptr, 1, 17, 0x0ull},
    {"topdown_l1", "BDW", 6, kNick18, 0, nullptr, nullptr, 1, 18, 0x4ull},
    {"topdown_l1", "SKX", 6, kNick19, 0, nullptr, nullptr, 1, 19, 0x4ull},
    {"topdown_l3_L1_bound", "BDX", 3, kNick20, 1, kRatio20, kRatioName20, 1, 20, 0x0ull},
    {"topdown_l3_L2_bound", "BDX", 3, kNick21, 0, nullptr, nullptr, 1, 21, 0x0ull},
    {"topdown_l3_icache", "BDX", 2, kNick22, 1, kRatio22, kRatioName22, 1, 22, 0x0ull},
    {"topdown_l4_mem", "BDX", 4, kNick23, 0, nullptr, nullptr, 1, 23, 0x0ull},
    {"llc_load_miss_rate", "any", 2, kNick24, 1, kRatio24, kRatioName24, 0, 24, 0x0ull},
    {"dtlb_load_miss_rate", "any", 2, kNick25, 1, kRatio25, kRatioName25, 0, 25, 0x0ull},
};

// the same ratio index pairs, [metric][ratio], for the device
__constant__ hbt_ratio c_metric_ratios[kNumMetrics * kMaxMetricRatios] = {
    {0, 0}, {0, 0},  // cpu_clock (any)
    {2, 1}, {1, 0},  // cs_ipc (any)
    {0, 0}, {0, 0},  // cycles (any)
    {1, 0}, {2, 0},  // cycles_breakdown (any)
    {1, 2}, {0, 0},  // dqos (any)
    {0, 0}, {0, 0},  // dram_access_reads (BDW)
    {0, 0}, {0, 0},  // dram_access_reads (BDW-DE)
    {0, 0}, {0, 0},  // dram_access_reads (BDX)
    {0, 0}, {0, 0},  // dram_access_reads (SKL)
    {0, 0}, {0, 0},  // dram_access_reads (SKX)
    {0, 0}, {0, 0},  // fp_instrs_double_precision (any)
    {0, 0}, {0, 0},  // fp_instrs_single_precision (any)
    {0, 0}, {0, 0},  // generic_sw (any)
    {0, 0}, {0, 0},  // instructions (any)
    {1, 0}, {0, 0},  // ipc (any)
    {0, 1}, {0, 0},  // l3_cache_misses_per_instruction (any)
    {0, 0}, {0, 0},  // page_faults (any)
    {0, 0}, {0, 0},  // system_calls (any)
    {0, 0}, {0, 0},  // topdown_l1 (BDW)
    {0, 0}, {0, 0},  // topdown_l1 (SKX)
    {0, 2}, {0, 0},  // topdown_l3_L1_bound (BDX)
    {0, 0}, {0, 0},  // topdown_l3_L2_bound (BDX)
    {0, 1}, {0, 0},  // topdown_l3_icache (BDX)
    {0, 0}, {0, 0},  // topdown_l4_mem (BDX)
    {0, 1}, {0, 0},  // llc_load_miss_rate (any)
    {0, 1}, {0, 0},  // dtlb_load_miss_rate (any)
};

}  // namespace hbt
