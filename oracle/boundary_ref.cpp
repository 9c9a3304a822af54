// TEST INFRASTRUCTURE -- not part of the product path.
//
// boundary_ref.cpp: the drop-in boundary exercised with the REFERENCE's own types (headers
// included from /root/reference, never copied).  Built by oracle/Makefile into
// oracle/_ref/boundary_ref (only where the reference tree is present; the binary travels to the
// GPU box) and run by tests/test_boundary_ref.py under -m gpu.
//
//  1. ring path: a real hbt::perf_event::CpuCountSampleGenerator per CPU over a synthetic
//     perf_event_mmap_page.  Its own onCpuDataBufferRead (CpuEventsGroup.h:1524-1570) feeds
//     hbt_submit_ring through hbt_b200::submitUnreadBytes (dynolog-1_b200/host/hbt_glue.h);
//     hbt_decode + hbt_fetch_results / hbt_fetch_rows / hbt_fetch_carry must then equal what THE
//     SAME OBJECT's consume() returns, counts and writes to its own output ring -- over two
//     consecutive batches, so the generator state carried on the GPU is compared with the
//     object's own members as well.
//  2. counter groups: hbt_counts_pipeline_host sums land in a real
//     GroupReadValues<mode::Counting> (hbt_b200::storeReadValues); its getCount() /
//     getTimeRunning() must equal the GPU's scaled counts and what the reference's own
//     PerCpuBase::read + getCount give on the same trace; the scaled counts then go through the
//     reference's unmodified dynolog::MetricSeries (diff) and must reproduce the GPU's deltas.
//
// Exit code 0 = every comparison held; each failure prints a line and counts.
#include <fcntl.h>
#include <linux/perf_event.h>
#include <unistd.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include <fmt/format.h>

#define protected public
#include "hbt/src/perf_event/PerCpuCountReader.h"
#include "hbt/src/perf_event/PerCpuCountSampleGenerator.h"
#include "hbt/src/ringbuffer/Consumer.h"
#undef protected
#include "dynolog/src/metric_frame/MetricSeries.h"

#include "../dynolog-1_b200/host/hbt_glue.h"
#include "../include/hbt_b200.h"

namespace hp = facebook::hbt::perf_event;
namespace hb = facebook::hbt;

static int g_failures = 0;
#define CHECK(cond, ...)                        \
  do {                                          \
    if (!(cond)) {                              \
      ++g_failures;                             \
      std::printf("FAIL %s:%d: ", __FILE__, __LINE__); \
      std::printf(__VA_ARGS__);                 \
      std::printf("\n");                        \
    }                                           \
  } while (0)

static uint64_t splitmix(uint64_t& s) {
  uint64_t z = (s += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

// A perf mmap area as the kernel lays it out: header page, then the data area.
struct FakePerfMmap {
  std::vector<uint8_t> mem;  // 4096 + data_size
  uint64_t data_size;
  explicit FakePerfMmap(uint64_t ds) : mem(4096 + ds, 0), data_size(ds) {
    auto* h = hdr();
    h->data_offset = 4096;
    h->data_size = ds;
    h->time_mult = 1;
    h->cap_user_time = 1;
    h->cap_user_time_zero = 1;
  }
  perf_event_mmap_page* hdr() { return reinterpret_cast<perf_event_mmap_page*>(mem.data()); }
  uint8_t* data() { return mem.data() + 4096; }
  // append words at data_head (the kernel's side of the ring)
  void produce(const std::vector<uint64_t>& words) {
    uint64_t head = hdr()->data_head;
    for (uint64_t w : words) {
      std::memcpy(data() + (head & (data_size - 1)), &w, 8);
      head += 8;
    }
    hdr()->data_head = head;
  }
};

static int devnull_fd() {
  static int fd = ::open("/dev/null", O_RDONLY);
  return fd;
}

template <class G>
static void attach(G& g, FakePerfMmap& mm) {
  g.mmap_base_ = mm.mem.data();
  g.event_fds_[0] = devnull_fd();  // isOpen() (CpuEventsGroup.h:581-583)
  g.tsc_rem_flags_ = 0;
}
template <class G>
static void detach(G& g) {  // keep ~CpuEventsGroupBase from munmap()ing the vector
  g.event_fds_[0] = -1;
  g.mmap_base_ = nullptr;
}

// mode::Sampling stream: SAMPLE = hdr tstamp nr time_enabled time_running value[nr]; now and
// then a LOST (hdr id num_lost) in front of one.
static std::vector<uint64_t> sampling_batch(uint64_t& seed, uint32_t n_samples, uint32_t E, uint64_t& ts,
                                            std::vector<uint64_t>& value, uint32_t lost_every) {
  std::vector<uint64_t> w;
  for (uint32_t k = 0; k < n_samples; ++k) {
    if (lost_every && k % lost_every == lost_every - 1) {
      w.push_back(PERF_RECORD_LOST | (24ull << 48));
      w.push_back(7);
      w.push_back(1 + splitmix(seed) % 50);
    }
    ts += 1000 + splitmix(seed) % 9000;
    w.push_back(PERF_RECORD_SAMPLE | ((uint64_t)(40 + 8 * E) << 48));
    w.push_back(ts);
    w.push_back(E);
    w.push_back(ts - 1000000);
    w.push_back(ts - 1000000 - splitmix(seed) % 7);
    for (uint32_t e = 0; e < E; ++e) {
      value[e] += 100000 + splitmix(seed) % 9900000;
      w.push_back(value[e]);
    }
  }
  return w;
}

static hp::EventConfs make_confs(uint32_t n_events) {
  hp::EventConfs confs;
  for (uint32_t i = 0; i < n_events; ++i) {
    hp::EventConf c;
    c.id = "ev" + std::to_string(i);
    c.configs = hp::EventConfigs{.type = 0, .config = i};
    confs.push_back(c);
  }
  return confs;
}

static void ring_path(hbt_ctx* ctx, uint32_t n_cpus, uint32_t E, uint64_t data_size, uint32_t lost_every,
                      uint32_t max_records) {
  using Gen = hp::CpuCountSampleGenerator;
  auto confs = make_confs(E);
  std::vector<std::string> names;
  for (auto& c : confs) names.push_back(c.id);
  const uint32_t per_batch = (uint32_t)((data_size / 2 - 4096) / (40 + 8 * E + 24));
  std::vector<std::unique_ptr<FakePerfMmap>> mms;
  std::vector<std::unique_ptr<Gen>> gens;
  std::vector<uint64_t> seeds(n_cpus), tss(n_cpus, 1000000);
  std::vector<std::vector<uint64_t>> values(n_cpus, std::vector<uint64_t>(E, 0));
  int rc = hbt_ring_configure(ctx, n_cpus, data_size);
  CHECK(rc == 0, "hbt_ring_configure %d", rc);
  const uint64_t seed_tstamp = 999000;  // what enableImpl() takes from now() (:91-106)
  hbt_ring_reset_state(ctx, seed_tstamp);
  for (uint32_t c = 0; c < n_cpus; ++c) {
    mms.push_back(std::make_unique<FakePerfMmap>(data_size));
    // start a few laps in, close to the ring end: the first window wraps
    const uint64_t start = 3 * data_size + data_size - 4096 - 8 * c;
    mms[c]->hdr()->data_head = mms[c]->hdr()->data_tail = start;
    gens.push_back(std::make_unique<Gen>(names, c, -1, confs, 2 * per_batch + 8, 0));
    attach(*gens[c], *mms[c]);
    gens[c]->count_.tstamp = seed_tstamp;  // the state enableImpl() leaves (:91-106), with a fixed now()
    for (uint32_t e = 0; e < E; ++e) gens[c]->last_[e] = 0;
    seeds[c] = 0x5EED000000000000ull + 7777 * E + c;
  }
  for (int batch = 0; batch < 2; ++batch) {
    for (uint32_t c = 0; c < n_cpus; ++c) {
      mms[c]->produce(sampling_batch(seeds[c], per_batch - 13 * c, E, tss[c], values[c], lost_every));
      // the reference's own onCpuDataBufferRead hands the unread bytes to the GPU library
      rc = hbt_b200::submitUnreadBytes(ctx, c, *gens[c], mms[c]->hdr());
      CHECK(rc == 0, "submitUnreadBytes cpu %u: %d (%s)", c, rc, hbt_last_error(ctx));
    }
    hbt_decode_desc d{};
    d.handled_mask = HBT_MODE_SAMPLING;
    d.sample_layout = HBT_SAMPLE_GROUP;
    d.aggregate = HBT_AGG_COUNT_DELTAS;
    d.n_events = E;
    d.max_records = max_records;
    d.tsc.time_mult = 1;
    rc = hbt_decode(ctx, &d);
    CHECK(rc == 0, "hbt_decode %d (%s)", rc, hbt_last_error(ctx));
    std::vector<hbt_cpu_result> res(n_cpus);
    rc = hbt_fetch_results(ctx, res.data(), n_cpus);
    CHECK(rc == (int)n_cpus, "hbt_fetch_results %d (%s)", rc, hbt_last_error(ctx));
    for (uint32_t c = 0; c < n_cpus; ++c) {
      Gen& g = *gens[c];
      // the same object's own consume(): return value, tail, counters, rows, state
      const ssize_t ret = g.consume(max_records);
      const uint64_t tail = mms[c]->hdr()->data_tail;
      CHECK((int64_t)ret == res[c].ret, "batch %d cpu %u: consume() returned %zd, GPU %lld", batch, c, ret,
            (long long)res[c].ret);
      CHECK(tail == res[c].new_tail, "batch %d cpu %u: data_tail %llu vs %llu", batch, c,
            (unsigned long long)tail, (unsigned long long)res[c].new_tail);
      CHECK(g.getNumRecordSample() == res[c].counters[HBT_CNT_SAMPLE] &&
                g.getNumRecordLost() == res[c].counters[HBT_CNT_LOST] &&
                g.getNumRecordRead() == res[c].counters[HBT_CNT_READ] &&
                g.getNumRecordThrottle() == res[c].counters[HBT_CNT_THROTTLE],
            "batch %d cpu %u: per-type counters differ", batch, c);
      CHECK(g.count_.tstamp == res[c].carry_tstamp, "batch %d cpu %u: count_.tstamp", batch, c);
      CHECK(g.num_unknown_count_interval_ == res[c].num_unknown_count_interval,
            "batch %d cpu %u: num_unknown_count_interval_ %u vs %llu", batch, c,
            (unsigned)g.num_unknown_count_interval_, (unsigned long long)res[c].num_unknown_count_interval);
      std::vector<uint64_t> want;
      {
        hb::ringbuffer::Consumer<Gen::RbExtraData> cons(g.getOutput());
        std::vector<uint64_t> row(2 + E);
        while (cons.copy(g.kWriteByteSize, row.data()) >= 0) want.insert(want.end(), row.begin(), row.end());
      }
      CHECK(want.size() == res[c].n_rows * (2 + E), "batch %d cpu %u: %zu rows in the generator's ring, GPU %llu",
            batch, c, want.size() / (2 + E), (unsigned long long)res[c].n_rows);
      std::vector<uint64_t> got(res[c].n_rows * (2 + E) + 1);
      const int64_t n = hbt_fetch_rows(ctx, c, got.data(), res[c].n_rows);
      CHECK(n == (int64_t)res[c].n_rows, "hbt_fetch_rows %lld", (long long)n);
      got.resize(res[c].n_rows * (2 + E));
      CHECK(got == want, "batch %d cpu %u: delta rows differ from the generator's output ring", batch, c);
      std::vector<uint64_t> last(E);
      hbt_fetch_carry(ctx, c, last.data(), E);
      bool same = true;
      for (uint32_t e = 0; e < E; ++e) same = same && last[e] == g.last_[e];
      CHECK(same, "batch %d cpu %u: last_[] differs", batch, c);
      // (the window the GPU holds advanced to new_tail, like the page's data_tail did)
    }
  }
  for (auto& g : gens) detach(*g);
  std::printf("ring path: %u CPUs x 2 batches, %u events, lost_every %u, max_records %u: %s\n", n_cpus, E,
              lost_every, max_records, g_failures ? "see failures" : "ok");
}

struct ReplayCpuReader {  // a CpuCountReader whose read() replays a recorded group read
  using TMode = hp::mode::Counting;
  const uint64_t* cur = nullptr;
  uint32_t n_events = 0;
  bool isOpen() const { return true; }
  bool isEnabled() const { return true; }
  void close() {}
  void enable() {}
  void disable() {}
  bool read(hp::GroupReadValues<TMode>& rv) const {
    if (cur[0] != n_events) return false;  // short read (CpuEventsGroup.h:596-615)
    std::memcpy(rv.t, cur, 8 * (3 + n_events));
    return true;
  }
};
struct ReplayPerCpu : public hp::PerCpuBase<ReplayCpuReader> {
  ReplayPerCpu(const hb::CpuSet& cpus, uint32_t n_events) : hp::PerCpuBase<ReplayCpuReader>(cpus, nullptr) {
    for_each_cpu(cpu, cpus) {
      cpu_generators_[cpu] = std::make_shared<ReplayCpuReader>();
      cpu_generators_[cpu]->n_events = n_events;
    }
  }
};

static void counting_path(hbt_ctx* ctx, uint64_t I, uint32_t C, uint32_t E) {
  using RV = hp::GroupReadValues<hp::mode::Counting>;
  const uint64_t W = 3 + E;
  std::vector<uint64_t> reads(I * C * W);
  uint64_t seed = 424242 + E;
  std::vector<uint64_t> te(C, 0), tr(C, 0), cnt((size_t)C * E, 0);
  for (uint64_t i = 0; i < I; ++i)
    for (uint32_t c = 0; c < C; ++c) {
      uint64_t* t = &reads[(i * C + c) * W];
      const uint64_t dte = 1000000000ull - 50000 + splitmix(seed) % 100000;
      te[c] += dte;
      tr[c] += dte / 4 + splitmix(seed) % (3 * dte / 4);  // multiplexed: 0 < running <= enabled
      t[0] = E;
      t[1] = te[c];
      t[2] = tr[c];
      for (uint32_t e = 0; e < E; ++e) t[3 + e] = (cnt[c * E + e] += 100000 + splitmix(seed) % 9900000);
    }
  reads[(I / 2 * C + 1) * W] = 0;  // one failed read: that interval's PerCpuBase::read returns false
  std::vector<uint64_t> sums(I * W), scaled(I * E), deltas(I * E);
  std::vector<float> rates(I * E);
  int rc = hbt_counts_pipeline_host(ctx, reads.data(), I, C, E, nullptr, 0, sums.data(), scaled.data(),
                                    rates.data(), nullptr, deltas.data());
  CHECK(rc == 0, "hbt_counts_pipeline_host %d (%s)", rc, hbt_last_error(ctx));

  std::vector<unsigned> ids(C);
  for (uint32_t c = 0; c < C; ++c) ids[c] = c;
  hb::CpuSet cpus(hb::CpuSet::make_cpu_set_t(ids), C - 1, C - 1);
  ReplayPerCpu per_cpu(cpus, E);
  RV rv_ref(E), rv_gpu(E);
  // the reference's metric table: one series of scaled counts per event (MetricSeries.h)
  std::vector<facebook::dynolog::MetricSeries<uint64_t>> series;
  for (uint32_t e = 0; e < E; ++e) series.emplace_back(2, "ev" + std::to_string(e), "scaled count");
  for (uint64_t i = 0; i < I; ++i) {
    for (uint32_t c = 0; c < C; ++c) per_cpu.cpu_generators_[c]->cur = &reads[(i * C + c) * W];
    const bool ok_ref = per_cpu.read(rv_ref);                            // the reference's own reduce
    const bool ok_gpu = hbt_b200::storeReadValues(rv_gpu, &sums[i * W]);  // GPU sums in a real GroupReadValues
    CHECK(ok_ref == ok_gpu, "interval %llu: read ok %d vs %d", (unsigned long long)i, ok_ref, ok_gpu);
    for (uint32_t e = 0; e < E; ++e) {
      uint64_t want = 0;
      if (ok_ref) {
        CHECK(rv_gpu.getTimeRunning() == rv_ref.getTimeRunning() && rv_gpu.getTimeEnabled() == rv_ref.getTimeEnabled() &&
                  rv_gpu.getRawCount(e) == rv_ref.getRawCount(e),
              "interval %llu event %u: summed triple differs", (unsigned long long)i, e);
        want = rv_ref.getCount(e);
        // the real class's scaling over the GPU's sums == the GPU's own scaling
        CHECK(rv_gpu.getCount(e) == scaled[i * E + e], "interval %llu event %u: getCount over GPU sums %llu, GPU scaled %llu",
              (unsigned long long)i, e, (unsigned long long)rv_gpu.getCount(e), (unsigned long long)scaled[i * E + e]);
        // PerfMonitor::log's rate expression (dynolog/src/PerfMonitor.cpp:56-57) over the real object
        const float rate = static_cast<float>(static_cast<double>(want) * 1e3 / static_cast<double>(rv_ref.getTimeRunning()));
        CHECK(std::memcmp(&rate, &rates[i * E + e], 4) == 0, "interval %llu event %u: rate", (unsigned long long)i, e);
      }
      CHECK(want == scaled[i * E + e], "interval %llu event %u: scaled %llu, reference %llu", (unsigned long long)i, e,
            (unsigned long long)scaled[i * E + e], (unsigned long long)want);
      // unmodified MetricSeries: the difference of the last two samples is the GPU's delta
      series[e].addSample(want);
      const uint64_t d = i == 0 ? want : series[e].diff();
      CHECK(d == deltas[i * E + e], "interval %llu event %u: MetricSeries::diff %llu, GPU delta %llu",
            (unsigned long long)i, e, (unsigned long long)d, (unsigned long long)deltas[i * E + e]);
    }
  }
  std::printf("counting path: %llu intervals x %u CPUs x %u events: %s\n", (unsigned long long)I, C, E,
              g_failures ? "see failures" : "ok");
}

int main() {
  hbt_ctx* ctx = nullptr;
  int rc = hbt_create(&ctx, 0, HBT_MAX_CPUS, HBT_MAX_EVENTS);
  if (rc != 0) {
    std::printf("hbt_create failed: %d (no usable sm_100 device; there is no CPU fallback)\n", rc);
    return 2;
  }
  // keep the reference's per-LOST warnings out of the test log
  ::freopen("/dev/null", "w", stderr);
  ring_path(ctx, 4, 8, 1 << 20, 50, 0xFFFFFFFFu);
  ring_path(ctx, 3, 32, 1 << 21, 0, 0xFFFFFFFFu);
  ring_path(ctx, 2, 1, 1 << 18, 7, 1234);   // consume(max_num_records) stops inside the window
  counting_path(ctx, 400, 64, 8);
  counting_path(ctx, 50, 7, 40);
  hbt_destroy(ctx);
  std::printf("%s (%d failures)\n", g_failures ? "BOUNDARY MISMATCH" : "boundary ok", g_failures);
  return g_failures ? 1 : 0;
}
