// Tests of the host-side mirror (hbt_gpu.h, perf_monitor.h).  `host_test cpu` needs no
// GPU; `host_test gpu` drives the CUDA path through the mirrored reference API and
// checks it against the reference's expressions restated here (test-side only).
#include <dlfcn.h>

#include <cmath>
#include <cstdio>
#include <iostream>
#include <sstream>

#include "../perf_monitor.h"

using namespace hbt_gpu;
namespace pe = hbt_gpu::perf_event;

static int g_fail = 0;
#define CHECK(c)                                                          \
  do {                                                                    \
    if (!(c)) {                                                           \
      std::printf("FAIL %s:%d: %s\n", __FILE__, __LINE__, #c);            \
      ++g_fail;                                                           \
    }                                                                     \
  } while (0)
template <class E, class F>
bool throws(F f) {
  try {
    f();
  } catch (const E&) {
    return true;
  } catch (...) {
    return false;
  }
  return false;
}

static uint64_t sm64(uint64_t& s) {
  uint64_t z = (s += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

// reads[I][C][3+E], cumulative, multiplexed
static std::vector<uint64_t> make_trace(size_t I, size_t C, size_t E, uint64_t seed) {
  std::vector<uint64_t> r(I * C * (3 + E));
  std::vector<uint64_t> te(C, 0), tr(C, 0), cnt(C * E, 0);
  for (size_t i = 0; i < I; ++i)
    for (size_t c = 0; c < C; ++c) {
      uint64_t d = 1000000000ull + sm64(seed) % 100000;
      te[c] += d;
      tr[c] += d / 4 + sm64(seed) % (d - d / 4);
      uint64_t* t = &r[(i * C + c) * (3 + E)];
      t[0] = E;
      t[1] = te[c];
      t[2] = tr[c];
      for (size_t e = 0; e < E; ++e) {
        cnt[c * E + e] += 100000 + sm64(seed) % 9900000;
        t[3 + e] = cnt[c * E + e];
      }
    }
  return r;
}

static void test_cpu() {
  CHECK(throws<std::invalid_argument>([] { CpuSet::makeFromCont(std::vector<int>{1, 2, 600}); }));
  CHECK(throws<std::invalid_argument>([] { pe::GroupReadValues rv(0); }));
  pe::GroupReadValues rv(3);
  CHECK(rv.getNumEvents() == 3 && rv.getNumReadBytes() == 48);
  CHECK(throws<std::invalid_argument>([&] { rv.getRawCount(3); }));
  CHECK(!rv.hasMultiplexing());
  CHECK(rv.getRunningRatio() == 1.0f && rv.getRunningRatioInteger<>() == 100);  // time_enabled == 0
  rv.raw()[1] = 4000;
  rv.raw()[2] = 1000;
  CHECK(rv.hasMultiplexing() && rv.getRunningRatio() == 4.0f && rv.getRunningRatioInteger<1000>() == 250);
  std::ostringstream os;
  dynolog_gpu::JsonLogger lg(os);
  lg.logFloat("mips", 691.4972f);
  lg.logUint("cycles", 123456789012345ull);
  lg.logInt("neg", -5);
  lg.logStr("s", "x");
  CHECK(lg.sample().at("mips") == "\"691.497\"");  // "{:.3f}" as a string, Logger.cpp:42-44
  lg.finalize();
  CHECK(os.str() == "{\"cycles\":123456789012345,\"mips\":\"691.497\",\"neg\":-5,\"s\":\"x\"}\n");
  CHECK(lg.sample().empty());
  pe::MetricDesc md{"ipc", {{"cycles"}, {"instructions"}}};
  CHECK(*md.getNumEvents() == 2 && md.eventNicknames()[1] == "instructions");
  pe::MetricDesc none{"none", {}};
  CHECK(!none.getEventRefs().has_value());
  // no GPU in this process -> constructing the context must fail loudly
  int n = 0;
  bool have_gpu = false;
  {
    hbt_ctx* c = nullptr;
    int rc = hbt_create(&c, 0, 8, 8);
    have_gpu = rc == 0;
    if (c) hbt_destroy(c);
    (void)n;
  }
  if (!have_gpu) CHECK(throws<std::system_error>([] { GpuContext g(0); }));
  CHECK(throws<std::invalid_argument>([] { GpuContext g(0, HBT_MAX_CPUS + 1, 8); }));
}

// Row sink, CountData and the TraceMonitor consumers (no GPU needed).  When the reference
// harness (oracle/_ref/libhbt_ref.so, env HBT_REF_LIB) is present the sink is also checked
// against the reference's own ringbuffer::Producer/Consumer on random row sequences.
static void test_row_sink() {
  const size_t E = 3;
  auto row = [&](uint64_t k) { return std::vector<uint64_t>{1000 + 10 * k, 10, k, 2 * k, 3 * k}; };
  // 40-byte rows, 6 requested -> 256-byte pool: 6 rows fit, the 7th does not
  pe::CountSampleSink s0(E, 6, 0);
  CHECK(s0.kWriteByteSize == 40 && s0.kDataPoolByteSize == 256);
  for (uint64_t k = 0; k < 6; ++k) CHECK(s0.write(row(k).data()) == 0);
  CHECK(s0.write(row(6).data()) == -ENOSPC && s0.getNumEntries() == 6);
  // drop rule: the 2 oldest rows go, the new one enters
  pe::CountSampleSink s1(E, 6, 2);
  for (uint64_t k = 0; k < 7; ++k) CHECK(s1.write(row(k).data()) == 0);
  CHECK(s1.getNumEntries() == 5 && s1.peek(0)[0] == 1020 && s1.peek(4)[0] == 1060);
  CHECK(throws<std::invalid_argument>([] { pe::CountSampleSink bad(3, 2, 100); }));

  mon::CountData cd({"a", "b", "c"});
  CHECK(cd.kRowStride == 5 && !cd.getTimeStampRange().has_value());
  for (uint64_t k = 0; k < 4; ++k) cd.write(row(k).data());
  CHECK(cd.numIntervals() == 4);
  CHECK(cd.getTimeStampRange()->first == 1000 && cd.getTimeStampRange()->second == 1040);
  CHECK(*cd.findCountAt(1019) == 1 && *cd.findCountAt(1020) == 2 && !cd.findCountAt(1040).has_value());
  auto f = cd.filter({true, false, false, true});
  CHECK(f->numIntervals() == 2 && f->data[5] == 1030);

  auto cpus = CpuSet::makeFromCont(std::vector<int>{1, 4});
  mon::TraceMonitor tm(cpus, {"a", "b", "c"}, 64, 0);
  for (uint64_t k = 0; k < 5; ++k) tm.sink(4).write(row(k).data());
  for (uint64_t k = 0; k < 2; ++k) tm.sink(1).write(row(k + 10).data());
  CHECK(throws<std::invalid_argument>([&] { tm.sink(2); }));
  // rows ending before stop_ts are released, the one reaching it stays (TraceMonitor.cpp:62-70)
  auto rel = tm.releaseCountData(4, 1030);
  CHECK(rel->numIntervals() == 2 && tm.sink(4).getNumEntries() == 3);
  std::vector<std::vector<uint64_t>> seen;
  auto func = [&](CpuId cpu, int idx, const std::vector<std::string>& names, TimeStamp t, TimeStamp d,
                  const uint64_t* c, size_t n) {
    CHECK(names.size() == 3 && n == 3 && d == 10);
    seen.push_back({cpu, (uint64_t)idx, t, c[0], c[2]});
  };
  tm.applyToCountSamplesAndConsume(func);
  CHECK(seen.size() == 5);
  if (seen.size() == 5) {
    CHECK((seen[0] == std::vector<uint64_t>{1, 0, 1100, 10, 30}));  // CPUs ascending, idx per CPU
    CHECK((seen[2] == std::vector<uint64_t>{4, 0, 1020, 2, 6}));
    CHECK((seen[4] == std::vector<uint64_t>{4, 2, 1040, 4, 12}));
  }
  CHECK(tm.sink(1).getNumEntries() == 0 && tm.sink(4).getNumEntries() == 0);

  const char* ref = std::getenv("HBT_REF_LIB");
  void* h = ref ? dlopen(ref, RTLD_NOW | RTLD_LOCAL) : nullptr;
  using Fn = int (*)(const uint64_t*, uint64_t, uint32_t, uint64_t, uint64_t, uint64_t*, uint64_t, uint64_t*, uint64_t*);
  Fn fn = h ? (Fn)dlsym(h, "hbtref_row_sink") : nullptr;
  if (!fn) {
    std::printf("row sink vs reference ring buffer: skipped (HBT_REF_LIB not set or not loadable)\n");
    return;
  }
  uint64_t seed = 42;
  int cases = 0;
  for (uint32_t ev : {1u, 3u, 8u, 32u})
    for (uint64_t cap : {1ull, 5ull, 16ull, 100ull})
      for (uint64_t drop : {0ull, 1ull, 3ull}) {
        const uint64_t n = 1 + sm64(seed) % 300;
        std::vector<uint64_t> rows(n * (2 + ev));
        for (auto& w : rows) w = sm64(seed);
        std::vector<uint64_t> out(rows.size());
        uint64_t n_out = 0, refused = 0;
        const size_t pool = [&] { size_t p = 1; while (p < cap * 8 * (2 + ev)) p <<= 1; return p; }();
        if (drop * 8 * (2 + ev) > pool) continue;
        CHECK(fn(rows.data(), n, ev, cap, drop, out.data(), n, &n_out, &refused) == 0);
        pe::CountSampleSink sk(ev, cap, drop);
        const size_t my_refused = sk.write(rows.data(), n);
        CHECK(my_refused == refused && sk.getNumEntries() == n_out);
        bool same = sk.getNumEntries() == n_out;
        for (uint64_t r = 0; same && r < n_out; ++r)
          same = std::memcmp(sk.peek(r), &out[r * (2 + ev)], 8 * (2 + ev)) == 0;
        CHECK(same);
        ++cases;
      }
  std::printf("row sink vs reference ring buffer: %d cases\n", cases);
}

// hand-written mode::ContextSwitch stream (CpuEventsGroup.h:105-184)
static std::vector<uint64_t> switch_stream() {
  auto hdr = [](uint64_t t, uint64_t words, uint64_t misc = 0) { return t | (misc << 32) | ((8 * words) << 48); };
  const uint64_t OUT = 1u << 13, PRE = (1u << 13) | (1u << 14);
  std::vector<uint64_t> w;
  auto add = [&](std::initializer_list<uint64_t> l) { w.insert(w.end(), l); };
  add({hdr(15, 4), 0, 40ull | (41ull << 32), 100});                       // switch in tid 41
  add({hdr(15, 4, OUT), 0, 40ull | (41ull << 32), 110});                  // out, yield
  add({hdr(7, 6), 50ull | (1ull << 32), 51ull | (1ull << 32), 120, 0, 121});  // FORK tid 51
  add({hdr(3, 5), 50ull | (51ull << 32), 0x6161, 50ull | (51ull << 32), 130});  // COMM tid 51
  add({hdr(15, 4), 0, 50ull | (51ull << 32), 140});                       // switch in tid 51
  add({hdr(2, 5), 7, 3, 0, 150});                                         // LOST(3)
  add({hdr(15, 4, PRE), 0, 50ull | (51ull << 32), 170});                  // out, preempt
  add({hdr(15, 4), 0, 40ull | (41ull << 32), 180});                       // switch in tid 41 again
  add({hdr(4, 6), 50ull | (1ull << 32), 51ull | (1ull << 32), 200, 0, 201});  // EXIT tid 51
  return w;
}

static void test_vid_mapper() {
  using E = pe::Event;
  std::vector<E> ev = {
      {100, HBT_SWEV_SWITCH_IN, 2, 0, 41},           {110, HBT_SWEV_SWITCH_OUT_YIELD, 2, 0, HBT_TAG_NA},
      {120, HBT_SWEV_THREAD_CREATION, 2, 0, 51},     {130, HBT_SWEV_COMM, 2, 0, 51},
      {135, HBT_SWEV_COMM, 2, 0, 77},                {140, HBT_SWEV_SWITCH_IN, 2, 0, 0},
      {141, HBT_SWEV_WRITE_ERRORS_START, 0xFF, 0, HBT_TAG_NA}, {150, HBT_SWEV_WRITE_ERRORS_END, 0xFF, 0, HBT_TAG_NA},
      {150, HBT_SWEV_SWITCH_IN, 2, 0, 41},           {160, HBT_SWEV_THREAD_DESTRUCTION, 2, 0, 41},
      {170, HBT_SWEV_SWITCH_IN, 2, 0, 41}};
  pe::ThreadSwitchVidMapper m;
  CHECK(m.apply(ev) == 9);  // the two COMM pseudo-events are gone
  const uint32_t want[9] = {1, HBT_TAG_NA, 2, 0, HBT_TAG_NA, HBT_TAG_NA, 4, 4, 5};
  // tid 41 -> 1, tid 51 -> 2, COMM of unseen tid 77 takes vid 3, tid 0 stays 0,
  // the LOST forgets everybody: 41 -> 4; after its EXIT the same tid is a new thread: 5
  for (size_t i = 0; i < 9 && ev.size() == 9; ++i) CHECK(ev[i].tag == want[i]);
  CHECK(m.next_vid == 6 && m.active_tids.size() == 1);
}

static void test_gpu() {
  auto gpu = std::make_shared<GpuContext>(0);
  const size_t C = 16, I = 6;
  auto cpus = CpuSet::makeRange(C);
  auto metrics = std::make_shared<dynolog_gpu::Metrics>();
  metrics->descs["instructions"] = std::make_shared<pe::MetricDesc>(pe::MetricDesc{"instructions", {{"instructions"}}});
  metrics->descs["cycles"] = std::make_shared<pe::MetricDesc>(pe::MetricDesc{"cycles", {{"cycles"}}});
  metrics->descs["ipc"] =
      std::make_shared<pe::MetricDesc>(pe::MetricDesc{"ipc", {{"cycles"}, {"instructions"}}});
  auto tr_i = make_trace(I, C, 1, 11), tr_c = make_trace(I, C, 1, 22), tr_ipc = make_trace(I, C, 2, 33);
  tr_ipc[(3 * C + 5) * 5] = 0;  // interval 3, cpu 5: failed read -> nullopt for "ipc"
  std::map<std::string, std::shared_ptr<pe::CounterSource>> src;
  src["instructions"] = std::make_shared<pe::TraceReplaySource>(tr_i, C, 1);
  src["cycles"] = std::make_shared<pe::TraceReplaySource>(tr_c, C, 1);
  src["ipc"] = std::make_shared<pe::TraceReplaySource>(tr_ipc, C, 2);
  dynolog_gpu::PerfMonitor pm(cpus, {"instructions", "cycles", "ipc"}, metrics, src, gpu);
  for (size_t i = 0; i < I; ++i) {
    pm.step();
    std::ostringstream os;
    dynolog_gpu::JsonLogger lg(os);
    pm.log(lg);
    // expectations: the reference's expressions (PerCpuBase.h:104-119, CpuEventsGroup.h:449-459,
    // PerfMonitor.cpp:52-67) restated on the host
    auto expect = [&](const std::vector<uint64_t>& tr, size_t E, size_t e, uint64_t& time) {
      uint64_t te = 0, trn = 0, cnt = 0;
      for (size_t c = 0; c < C; ++c) {
        const uint64_t* t = &tr[(i * C + c) * (3 + E)];
        te += t[1];
        trn += t[2];
        cnt += t[3 + e];
      }
      time = trn;
      return te == 0 ? 0 : (uint64_t)((double)cnt * (double)te / (double)trn);
    };
    uint64_t time = 0;
    char buf[64];
    uint64_t ci = expect(tr_i, 1, 0, time);
    std::snprintf(buf, sizeof buf, "\"%.3f\"", (float)((double)ci * 1e3 / (double)time));
    CHECK(lg.sample().at("mips") == buf);
    uint64_t cc = expect(tr_c, 1, 0, time);
    std::snprintf(buf, sizeof buf, "\"%.3f\"", (float)((double)cc * 1e3 / (double)time));
    CHECK(lg.sample().at("mega_cycles_per_second") == buf);
    CHECK(pm.readValues().at("instructions")->getCount(0) == ci);
    if (i == 3) {
      CHECK(!pm.readValues().at("ipc").has_value());  // one CPU failed -> whole read fails
      CHECK(lg.sample().count("cycles") == 0);
    } else {
      // metric "ipc": neither event matches the (id, nickname) special cases -> logUint(nickname)
      uint64_t c0 = expect(tr_ipc, 2, 0, time), c1 = expect(tr_ipc, 2, 1, time);
      CHECK(lg.sample().at("cycles") == std::to_string(c0));
      CHECK(lg.sample().at("instructions") == std::to_string(c1));
      CHECK(pm.readValues().at("ipc")->hasMultiplexing());
    }
  }
  // Monitor: duplicate key, state machine, multiplexing queue
  mon::Monitor<> m;
  auto md = metrics->descs["cycles"];
  auto s2 = std::make_shared<pe::TraceReplaySource>(tr_c, C, 1);
  auto r1 = m.emplaceCountReader(std::string("g1"), "a", md, s2, cpus, gpu);
  auto r2 = m.emplaceCountReader(std::string("g2"), "b", md, s2, cpus, gpu);
  CHECK(throws<std::invalid_argument>([&] { m.emplaceCountReader(std::nullopt, "a", md, s2, cpus, gpu); }));
  CHECK(!r1->isOpen() && !m.isOpen());
  CHECK(m.open() && !m.open() && m.isOpen() && r1->isOpen() && r2->isOpen() && !r1->isEnabled());
  CHECK(m.enable() && m.isEnabled() && !m.isOpen());
  CHECK(r1->isEnabled() && !r2->isEnabled());  // only the front mux group runs
  m.muxRotate();
  CHECK(!r1->isEnabled() && r2->isEnabled());
  CHECK(m.disable() && !r2->isEnabled() && r2->isOpen());
  CHECK(m.close() && !r1->isOpen());
  CHECK(!m.readCounts("a").has_value());  // closed reader: read() fails
  CHECK(throws<std::invalid_argument>([&] { r1->enable(); }));

  // ring decode through the mirrored generator API: LOST(num_lost=3) + 5 SAMPLEs of 2 events
  const size_t ds = 1 << 16;
  std::vector<uint64_t> ring(ds / 8, 0);
  const uint64_t tail = 5 * ds + ds - 64;  // wraps the ring end
  std::vector<uint64_t> s;
  s.insert(s.end(), {2ull | (24ull << 48), 7, 3});
  for (uint64_t k = 0; k < 5; ++k) s.insert(s.end(), {9ull | (56ull << 48), 1000 + 10 * k, 2, 5, 5, 100 * k, 200 * k});
  for (size_t w = 0; w < s.size(); ++w) ring[((tail / 8) + w) & (ds / 8 - 1)] = s[w];
  const uint64_t head = tail + 8 * s.size();
  pe::PerCpuCountSampleDecoder dec(CpuSet::makeFromCont(std::vector<int>{0, 2}), 2, ds, gpu);
  dec.enable(kInvalidTimeStamp);
  const uint8_t* base = reinterpret_cast<const uint8_t*>(ring.data());
  // the two slices onCpuDataBufferRead reports for a wrapped window; the second over-reports
  pe::RbDataSlices slices(base + (tail & (ds - 1)), 64, base, ds - 64);
  dec.onCpuDataBufferRead(2, slices, tail, head);
  auto ret = dec.consume(0xFFFFFFFFu);
  CHECK(ret.at(2) == -ENODATA && ret.at(0) == -ENODATA);
  CHECK(dec.getNumRecordLost(2) == 3 && dec.getNumRecordSample(2) == 5);
  CHECK(dec.getNumUnknownCountInterval(2) == 1 && dec.dataTail(2) == head);
  auto rows = dec.rows(2);
  CHECK(rows.size() == 4 * 4);
  for (uint64_t k = 0; k < 4 && rows.size() == 16; ++k) {
    CHECK(rows[4 * k] == 1000 + 10 * k && rows[4 * k + 1] == 10);
    CHECK(rows[4 * k + 2] == 100 && rows[4 * k + 3] == 200);
  }
  CHECK(dec.getLastTimeStamp(2) == 1040);
  CHECK(throws<std::invalid_argument>([&] { dec.onCpuDataBufferRead(1, slices, tail, head); }));
  // rows -> sinks -> applyToCountSamplesAndConsume callback (TraceMonitor.h:45-109)
  mon::TraceMonitor tm(CpuSet::makeFromCont(std::vector<int>{0, 2}), {"cycles", "instructions"}, 16, 0);
  CHECK(tm.accum(dec) == 0 && tm.sink(2).getNumEntries() == 4 && tm.sink(0).getNumEntries() == 0);
  int calls = 0;
  auto func = [&](CpuId cpu, int idx, const std::vector<std::string>& names, TimeStamp t, TimeStamp d,
                  const uint64_t* c, size_t n) {
    CHECK(cpu == 2 && idx == calls && names[1] == "instructions" && n == 2);
    CHECK(t == 1000 + 10 * (uint64_t)idx && d == 10 && c[0] == 100 && c[1] == 200);
    ++calls;
  };
  tm.applyToCountSamplesAndConsume(func);
  CHECK(calls == 4 && tm.sink(2).getNumEntries() == 0);

  // back-pressure: a sink with room for 2 rows stops consume() in front of the SAMPLE whose row
  // would be the third (-ENOSPC, tail before it); once drained, the rest follows and that SAMPLE re-seeds
  {
    mon::TraceMonitor small(CpuSet::makeFromCont(std::vector<int>{0, 2}), {"cycles", "instructions"}, 2, 0);
    const size_t room = small.sink(2).freeRows();
    dec.enable(kInvalidTimeStamp);
    dec.onCpuDataBufferRead(2, slices, tail, head);
    small.applyBackPressure(dec);
    auto r1 = dec.consume(0xFFFFFFFFu);
    CHECK(r1.at(2) == -ENOSPC && dec.rows(2).size() == room * 4);
    CHECK(small.accum(dec) == 0 && small.sink(2).getNumEntries() == room && small.sink(2).freeRows() == 0);
    // LOST + re-seeding SAMPLE + `room` row-writing SAMPLEs were consumed; the tail stands in front of the next
    CHECK(dec.dataTail(2) == tail + 24 + 56 * (1 + room));
    CHECK(dec.getLastTimeStamp(2) == kInvalidTimeStamp);
    int seen = 0;
    auto drain = [&](CpuId, int, const std::vector<std::string>&, TimeStamp, TimeStamp, const uint64_t*, size_t) { ++seen; };
    small.applyToCountSamplesAndConsume(drain);
    CHECK(seen == (int)room);
    // the rest of the window (what onCpuDataBufferRead hands out from the new tail; it no longer wraps)
    const uint64_t t2 = dec.dataTail(2);
    pe::RbDataSlices rest(base + (t2 & (ds - 1)), head - t2);
    dec.onCpuDataBufferRead(2, rest, t2, head);
    small.applyBackPressure(dec);
    auto r2 = dec.consume(0xFFFFFFFFu);
    CHECK(r2.at(2) == -ENODATA && dec.dataTail(2) == head);
    CHECK(dec.rows(2).size() == (4 - room - 1) * 4);   // the refused SAMPLE only re-seeds
    dec.clearRowsBudget();
  }

  // context-switch ring through the thread-switch generator mirror: events with vids
  {
    const std::vector<uint64_t> sw = switch_stream();
    std::vector<uint64_t> sring(ds / 8, 0);
    const uint64_t stail = 3 * ds + 40;
    for (size_t w = 0; w < sw.size(); ++w) sring[((stail / 8) + w) & (ds / 8 - 1)] = sw[w];
    const uint64_t shead = stail + 8 * sw.size();
    pe::PerCpuThreadSwitchDecoder sdec(CpuSet::makeFromCont(std::vector<int>{1, 3}), 2, ds, gpu);
    const uint8_t* sbase = reinterpret_cast<const uint8_t*>(sring.data());
    pe::RbDataSlices ss(sbase + (stail & (ds - 1)), 8 * sw.size());
    sdec.onCpuDataBufferRead(3, ss, stail, shead);
    auto sret = sdec.consume(0xFFFFFFFFu);
    CHECK(sret.at(3) == -ENODATA && sret.at(1) == -ENODATA);
    const auto& ev = sdec.events(3);
    struct W { uint64_t ts; uint8_t type; uint32_t tag; };
    const W want[] = {{100, HBT_SWEV_SWITCH_IN, 1},          {110, HBT_SWEV_SWITCH_OUT_YIELD, HBT_TAG_NA},
                      {120, HBT_SWEV_THREAD_CREATION, 2},    {140, HBT_SWEV_SWITCH_IN, 2},
                      {141, HBT_SWEV_WRITE_ERRORS_START, HBT_TAG_NA}, {170, HBT_SWEV_WRITE_ERRORS_END, HBT_TAG_NA},
                      {170, HBT_SWEV_SWITCH_OUT_PREEMPT, HBT_TAG_NA}, {180, HBT_SWEV_SWITCH_IN, 3},
                      {200, HBT_SWEV_THREAD_DESTRUCTION, 4}};
    CHECK(ev.size() == 9);
    for (size_t i = 0; i < 9 && ev.size() == 9; ++i) {
      CHECK(ev[i].tstamp == want[i].ts && ev[i].type == want[i].type && ev[i].tag == want[i].tag);
      CHECK(ev[i].comp_unit_id == 3);
      CHECK(ev[i].level == (ev[i].type >= HBT_SWEV_WRITE_ERRORS_START ? HBT_LEVEL_INVALID : 2));
    }
    CHECK(sdec.getNumRecordLost(3) == 3 && sdec.getNumRecordSwitchCpuWide(3) == 5);
    CHECK(sdec.getLastTimeStamp(3) == 200 && sdec.dataTail(3) == shead && sdec.events(1).empty());
  }
}

int main(int argc, char** argv) {
  const std::string mode = argc > 1 ? argv[1] : "cpu";
  test_cpu();
  test_row_sink();
  test_vid_mapper();
  if (mode == "gpu") test_gpu();
  std::printf("%s: %s (%d failures)\n", mode.c_str(), g_fail ? "FAILED" : "ok", g_fail);
  return g_fail ? 1 : 0;
}
