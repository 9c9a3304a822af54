/*
 * TEST INFRASTRUCTURE -- not part of the product path.
 *
 * hbt_oracle.c: plain-C, single-threaded restatement ("port") of the reference
 * algorithm for the perf_event decode/aggregate hot path.  Every function cites
 * the reference file:line it follows (paths under the reference tree).
 *
 * Pinning: oracle/_ref/libhbt_ref.so drives the reference's own headers
 * (oracle/ref_harness.cpp); tests/test_oracle_vs_reference.py checks this file
 * against it on seeded inputs, and tests/golden/ holds vectors generated from
 * it by tests/golden/make_golden.py, so this file is checked even where the
 * reference tree is absent.  Two parts have NO reference implementation and are
 * therefore "parity unpinned" beyond the consume() loop around them: the
 * (cpu,pid,ip) histogram (hbt_oracle_ip_hist_*) and the derived ratios.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load
 * this library.  The product (libhbt_b200.so) never links or calls it.
 */
#include <errno.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/hbt_b200.h"

/* PERF_RECORD_* values, /usr/include/linux/perf_event.h:846-1260. */
enum {
  REC_LOST = 2,
  REC_COMM = 3,
  REC_EXIT = 4,
  REC_THROTTLE = 5,
  REC_UNTHROTTLE = 6,
  REC_FORK = 7,
  REC_READ = 8,
  REC_SAMPLE = 9,
  REC_AUX = 11,
  REC_ITRACE_START = 12,
  REC_SWITCH_CPU_WIDE = 15
};

#define INVALID_TS UINT64_MAX /* kInvalidTimeStamp, hbt/src/common/System.h:97 */

int hbt_oracle_abi_version(void) { return HBT_B200_ABI_VERSION; }

/* CpuEventsGroup::kernelTimeFromTsc, CpuEventsGroup.h:1577-1591, with
 * tsc_rem_flags_ = (1 << time_shift) - 1 (:1146). */
uint64_t hbt_oracle_kernel_time_from_tsc(uint64_t tsc_cycles, const hbt_tsc_params* p) {
  uint64_t rem_flags = (((uint64_t)1) << p->time_shift) - 1;
  uint64_t quot = tsc_cycles >> p->time_shift;
  uint64_t rem = tsc_cycles & rem_flags;
  return p->time_zero + quot * p->time_mult + ((rem * p->time_mult) >> p->time_shift);
}

/* ---- (pid, ip) -> count open-addressing map for one CPU ------------------ */
typedef struct ip_slot {
  uint64_t ip;
  uint64_t count; /* 0 = empty */
  uint32_t pid;
} ip_slot;

typedef struct hbt_oracle_hist {
  ip_slot* slots;
  uint64_t cap; /* power of two */
  uint64_t size;
} hbt_oracle_hist;

static uint64_t mix64(uint64_t x) {
  x ^= x >> 33;
  x *= 0xff51afd7ed558ccdull;
  x ^= x >> 33;
  x *= 0xc4ceb9fe1a85ec53ull;
  x ^= x >> 33;
  return x;
}

hbt_oracle_hist* hbt_oracle_hist_new(void) {
  hbt_oracle_hist* h = (hbt_oracle_hist*)calloc(1, sizeof(*h));
  h->cap = 1024;
  h->slots = (ip_slot*)calloc(h->cap, sizeof(ip_slot));
  return h;
}

void hbt_oracle_hist_free(hbt_oracle_hist* h) {
  if (!h) return;
  free(h->slots);
  free(h);
}

static void hist_add(hbt_oracle_hist* h, uint64_t ip, uint32_t pid, uint64_t n);

static void hist_grow(hbt_oracle_hist* h) {
  ip_slot* old = h->slots;
  uint64_t old_cap = h->cap;
  h->cap *= 2;
  h->slots = (ip_slot*)calloc(h->cap, sizeof(ip_slot));
  h->size = 0;
  for (uint64_t i = 0; i < old_cap; ++i)
    if (old[i].count) hist_add(h, old[i].ip, old[i].pid, old[i].count);
  free(old);
}

static void hist_add(hbt_oracle_hist* h, uint64_t ip, uint32_t pid, uint64_t n) {
  if ((h->size + 1) * 2 > h->cap) hist_grow(h);
  uint64_t i = mix64(ip ^ ((uint64_t)pid << 32 | pid)) & (h->cap - 1);
  for (;;) {
    ip_slot* s = &h->slots[i];
    if (s->count == 0) {
      s->ip = ip;
      s->pid = pid;
      s->count = n;
      ++h->size;
      return;
    }
    if (s->ip == ip && s->pid == pid) {
      s->count += n;
      return;
    }
    i = (i + 1) & (h->cap - 1);
  }
}

uint64_t hbt_oracle_hist_size(const hbt_oracle_hist* h) { return h->size; }

uint64_t hbt_oracle_hist_fetch(const hbt_oracle_hist* h, uint32_t cpu, hbt_hist_entry* out,
                               uint64_t cap) {
  uint64_t n = 0;
  for (uint64_t i = 0; i < h->cap && n < cap; ++i) {
    if (h->slots[i].count) {
      out[n].ip = h->slots[i].ip;
      out[n].pid = h->slots[i].pid;
      out[n].cpu = cpu;
      out[n].count = h->slots[i].count;
      ++n;
    }
  }
  return n;
}

/* ---- consume ------------------------------------------------------------- */

/* State the reference keeps in the generator object across consume() calls:
 * CpuCountSampleGenerator::count_.tstamp, last_[], num_unknown_count_interval_
 * (PerCpuCountSampleGenerator.h:222-230). */
typedef struct hbt_oracle_carry {
  uint64_t tstamp;
  uint64_t last[HBT_MAX_EVENTS];
} hbt_oracle_carry;

static inline uint64_t ring_word(const uint8_t* ring, uint64_t mask, uint64_t pos) {
  /* Records are u64-aligned and the header never wraps (CpuEventsGroup.h:1353);
   * reading word by word through the mask is the reference's wrap linearisation
   * (:1358-1370) without the copy. */
  uint64_t w;
  memcpy(&w, ring + (pos & mask), 8);
  return w;
}

/* CpuEventsGroup::consume (CpuEventsGroup.h:1326-1522) for one ring, with the
 * handlers selected by desc:
 *   HBT_AGG_COUNT_DELTAS: CpuCountSampleGenerator::handleRecordLost / -Sample
 *                         (PerCpuCountSampleGenerator.h:110-124, :132-202)
 *   HBT_AGG_IP_HISTOGRAM: ++hist[(pid, ip)]   (no reference implementation)
 *   HBT_AGG_NONE:         handlers return 0   (CpuEventsGroupTest.cpp:58-133)
 * res->counters and res->num_unknown_count_interval are in/out running totals.
 * rows may be NULL unless COUNT_DELTAS; a full rows buffer behaves like the
 * reference's full output ring without drop-if-full (-ENOSPC, :191-194). */
/* PERF_RECORD_MISC_SWITCH_OUT / _PREEMPT, /usr/include/linux/perf_event.h:779-790 */
#define MISC_SWITCH_OUT (1u << 13)
#define MISC_SWITCH_OUT_PREEMPT (1u << 14)

/* One tagstack::Event (Event.h:28-60) as two words of the rows buffer. */
static int sw_write(uint64_t* rows, uint64_t rows_cap, uint64_t* n_rows, uint64_t tstamp,
                    uint32_t type, uint32_t level, uint32_t cu, uint32_t tag) {
  if (*n_rows >= rows_cap) return -ENOSPC; /* output_producer_.write fails */
  rows[2 * *n_rows] = tstamp;
  rows[2 * *n_rows + 1] = (uint64_t)(type & 0xFF) | ((uint64_t)(level & 0xFF) << 8) |
                          ((uint64_t)(cu & 0xFFFF) << 16) | ((uint64_t)tag << 32);
  ++*n_rows;
  return 0;
}

/* CpuThreadSwitchGenerator::pendingRecordLost_, PerCpuThreadSwitchGenerator.h:441-453.
 * carry->tstamp = last_tstamp_, carry->last[0] = in_write_errors_. */
static int sw_pending(hbt_oracle_carry* carry, uint64_t* rows, uint64_t rows_cap, uint64_t* n_rows,
                      uint64_t tstamp, uint32_t cu) {
  if (!carry->last[0]) return 0;
  int ret = sw_write(rows, rows_cap, n_rows, tstamp, HBT_SWEV_WRITE_ERRORS_END, HBT_LEVEL_INVALID, cu,
                     HBT_TAG_NA);
  if (ret >= 0) {
    carry->tstamp = tstamp;
    carry->last[0] = 0;
  }
  return ret;
}

static int consume_core(const uint8_t* ring, uint64_t data_size, uint64_t tail, uint64_t head,
                        const hbt_decode_desc* d, hbt_oracle_carry* carry, uint64_t* rows,
                        uint64_t rows_cap, hbt_oracle_hist* hist, hbt_cpu_result* res, uint32_t cpu) {
  const int sw = d->aggregate == HBT_AGG_SWITCH_EVENTS;
  int64_t err = 0;
  uint64_t n = 0;
  uint64_t n_rows = 0;
  const uint64_t mask = data_size - 1;
  const uint32_t hm = d->handled_mask;
  uint64_t* c = res->counters;

  res->n_rows = 0;
  if (data_size == 0) { /* :1342-1344 */
    res->ret = 0;
    res->n_records = 0;
    res->new_tail = tail;
    res->carry_tstamp = carry ? carry->tstamp : INVALID_TS;
    return 0;
  }

  while (tail < head && n < d->max_records) { /* :1348 */
    const uint64_t hdr = ring_word(ring, mask, tail);
    const uint32_t type = (uint32_t)hdr;
    const uint32_t size = (uint32_t)(hdr >> 48);
    if (size < 8 || (size & 7)) {
      /* Outside the kernel's guarantees (records are u64-padded).  The
       * reference never terminates on size 0; this port and the CUDA path stop
       * with -EBADMSG before the record. */
      err = -EBADMSG;
      goto exit;
    }
    switch (type) {
      case REC_LOST:
        if (hm & HBT_H_LOST) {
          /* Lost: hdr, id, num_lost (CpuEventsGroup.h:112-120, :236-242) */
          if (d->aggregate == HBT_AGG_COUNT_DELTAS) carry->tstamp = INVALID_TS; /* gen :122 */
          if (sw) {
            /* handleRecordLost, PerCpuThreadSwitchGenerator.h:228-259; ContextSwitch::Lost
             * = hdr id num_lost {pid tid tstamp} (CpuEventsGroup.h:112-120).  A record
             * shorter than its struct: the reference reads past it; here -EPERM. */
            if (size < 40) { err = -EPERM; goto exit; }
            const uint64_t ts = ring_word(ring, mask, tail + 32);
            int ret = sw_pending(carry, rows, rows_cap, &n_rows, ts, cpu);
            if (ret >= 0)
              ret = sw_write(rows, rows_cap, &n_rows, carry->tstamp + 1, HBT_SWEV_WRITE_ERRORS_START,
                             HBT_LEVEL_INVALID, cpu, HBT_TAG_NA);
            if (ret < 0) { err = ret; goto exit; } /* :1376-1379 */
            carry->tstamp = ts;
            carry->last[0] = 1;
          }
          c[HBT_CNT_LOST] += ring_word(ring, mask, tail + 16);                  /* :1381 */
        }
        break;
      case REC_COMM:
        if (hm & HBT_H_COMM) {
          if (sw) {
            /* handleRecordComm, :261-297: no Event, but the thread map changes -> pseudo-event.
             * Comm_Lower = hdr pid tid comm[]; sample_id closes the record (:1387-1393). */
            if (size < 32) { err = -EPERM; goto exit; }
            const uint64_t ts = ring_word(ring, mask, tail + size - 8);
            const uint32_t tid = (uint32_t)(ring_word(ring, mask, tail + 8) >> 32);
            int ret = sw_pending(carry, rows, rows_cap, &n_rows, ts, cpu);
            if (ret >= 0)
              ret = sw_write(rows, rows_cap, &n_rows, ts, HBT_SWEV_COMM, d->tid_level, cpu, tid);
            if (ret < 0) { err = ret; goto exit; }
            carry->tstamp = ts;
          }
          ++c[HBT_CNT_COMM]; /* :1394-1395 */
        }
        break;
      case REC_EXIT:
        if (hm & HBT_H_EXIT) {
          if (sw) {
            /* handleRecordExit, :299-331; Exit = hdr pid ppid tid ptid time sample_id (:139-146) */
            if (size < 48) { err = -EPERM; goto exit; }
            const uint64_t ts = ring_word(ring, mask, tail + 24);
            const uint32_t tid = (uint32_t)ring_word(ring, mask, tail + 16);
            int ret = sw_pending(carry, rows, rows_cap, &n_rows, ts, cpu);
            if (ret >= 0)
              ret = sw_write(rows, rows_cap, &n_rows, ts, HBT_SWEV_THREAD_DESTRUCTION, d->tid_level,
                             cpu, tid);
            if (ret < 0) { err = ret; goto exit; }
            carry->tstamp = ts;
          }
          ++c[HBT_CNT_EXIT];
        }
        break;
      case REC_THROTTLE:
        ++c[HBT_CNT_THROTTLE]; /* :1409-1414, unconditional */
        break;
      case REC_UNTHROTTLE:
        ++c[HBT_CNT_UNTHROTTLE];
        break;
      case REC_FORK:
        if (hm & HBT_H_FORK) {
          if (sw) { /* handleRecordFork, :333-373; Fork = Exit layout */
            if (size < 48) { err = -EPERM; goto exit; }
            const uint64_t ts = ring_word(ring, mask, tail + 24);
            const uint32_t tid = (uint32_t)ring_word(ring, mask, tail + 16);
            int ret = sw_pending(carry, rows, rows_cap, &n_rows, ts, cpu);
            if (ret >= 0)
              ret = sw_write(rows, rows_cap, &n_rows, ts, HBT_SWEV_THREAD_CREATION, d->tid_level, cpu,
                             tid);
            if (ret < 0) { err = ret; goto exit; }
            carry->tstamp = ts;
          }
          ++c[HBT_CNT_FORK];
        }
        break;
      case REC_SAMPLE:
        if (hm & HBT_H_SAMPLE) {
          if (d->sample_layout == HBT_SAMPLE_GROUP) {
            /* Sample: hdr tstamp nr time_enabled time_running value[nr]
             * (:268-281); recordSize = 40 + 8*nr (:330-333); check :1435-1440 */
            const uint64_t nr = ring_word(ring, mask, tail + 16);
            if ((uint64_t)size != 40 + 8 * nr) {
              err = -EPERM;
              goto exit;
            }
            if (d->aggregate == HBT_AGG_COUNT_DELTAS) {
              const uint64_t ts = ring_word(ring, mask, tail + 8);
              const uint32_t ne = d->n_events;
              if (carry->tstamp == INVALID_TS) { /* gen :135-154 */
                carry->tstamp = ts;
                for (uint32_t i = 0; i < ne; ++i) {
                  uint64_t v = ring_word(ring, mask, tail + 40 + 8ull * i);
                  if ((d->tsc_transform_mask >> i) & 1)
                    v = hbt_oracle_kernel_time_from_tsc(v, &d->tsc);
                  carry->last[i] = v;
                }
                ++res->num_unknown_count_interval;
              } else if (carry->tstamp > ts) { /* gen :156-163 */
                carry->tstamp = INVALID_TS;
                err = -EPERM;
                goto exit;
              } else {
                if (n_rows >= rows_cap) { /* write fails: gen :191-194 */
                  /* the deltas loop has already folded this sample into last_[] (:165-179) */
                  for (uint32_t i = 0; i < ne; ++i) {
                    uint64_t v = ring_word(ring, mask, tail + 40 + 8ull * i);
                    if ((d->tsc_transform_mask >> i) & 1)
                      v = hbt_oracle_kernel_time_from_tsc(v, &d->tsc);
                    carry->last[i] = v;
                  }
                  carry->tstamp = INVALID_TS;
                  err = -ENOSPC;
                  goto exit;
                }
                uint64_t* row = rows + n_rows * (2 + (uint64_t)ne);
                row[0] = carry->tstamp;      /* Count::tstamp, gen :223 */
                row[1] = ts - carry->tstamp; /* gen :165 */
                for (uint32_t i = 0; i < ne; ++i) {
                  uint64_t v = ring_word(ring, mask, tail + 40 + 8ull * i);
                  if ((d->tsc_transform_mask >> i) & 1)
                    v = hbt_oracle_kernel_time_from_tsc(v, &d->tsc);
                  row[2 + i] = v - carry->last[i]; /* gen :176 */
                  carry->last[i] = v;              /* gen :178 */
                }
                ++n_rows;
                carry->tstamp = ts; /* gen :199 */
              }
            }
          } else {
            /* IP layout: generic recordSize = sizeof(TRecord) */
            if (size != d->ip_sample_bytes) {
              err = -EPERM;
              goto exit;
            }
            if (d->aggregate == HBT_AGG_IP_HISTOGRAM) {
              const uint64_t ip = ring_word(ring, mask, tail + 8);
              const uint32_t pid = (uint32_t)ring_word(ring, mask, tail + 16);
              hist_add(hist, ip, pid, 1);
            }
          }
          ++c[HBT_CNT_SAMPLE]; /* :1445 */
        }
        break;
      case REC_READ:
        if (hm & HBT_H_READ) {
          /* Read: hdr pid tid nr time_enabled time_running value[nr]
           * (:254-265); check :1451-1456 */
          const uint64_t nr = ring_word(ring, mask, tail + 16);
          if ((uint64_t)size != 40 + 8 * nr) {
            err = -EPERM;
            goto exit;
          }
          ++c[HBT_CNT_READ];
        }
        break;
      case REC_AUX:
        if (hm & HBT_H_AUX) ++c[HBT_CNT_AUX];
        break;
      case REC_ITRACE_START:
        if (hm & HBT_H_ITRACE_START) ++c[HBT_CNT_ITRACE_START];
        break;
      case REC_SWITCH_CPU_WIDE:
        if (hm & HBT_H_SWITCH_CPU_WIDE) {
          if (sw) {
            /* handleRecordSwitchCpuWide, :375-412; SwitchCpuWide = hdr next_prev_pid
             * next_prev_tid {pid tid tstamp} (:162-183) */
            if (size < 32) { err = -EPERM; goto exit; }
            const uint64_t ts = ring_word(ring, mask, tail + 24);
            const uint32_t misc = (uint32_t)(hdr >> 32) & 0xFFFF;
            int ret = sw_pending(carry, rows, rows_cap, &n_rows, ts, cpu);
            if (ret >= 0) {
              if (misc & MISC_SWITCH_OUT)
                ret = sw_write(rows, rows_cap, &n_rows, ts,
                               (misc & MISC_SWITCH_OUT_PREEMPT) ? HBT_SWEV_SWITCH_OUT_PREEMPT
                                                                : HBT_SWEV_SWITCH_OUT_YIELD,
                               d->tid_level, cpu, HBT_TAG_NA);
              else
                ret = sw_write(rows, rows_cap, &n_rows, ts, HBT_SWEV_SWITCH_IN, d->tid_level, cpu,
                               (uint32_t)(ring_word(ring, mask, tail + 16) >> 32));
            }
            if (ret < 0) { err = ret; goto exit; }
            carry->tstamp = ts;
          }
          ++c[HBT_CNT_SWITCH_CPU_WIDE];
        }
        break;
      default: /* :1498-1500: logged, skipped by size */
        break;
    }
    ++n; /* :1502-1503 */
    tail += size;
  }
  if (tail == head && n < d->max_records) err = -ENODATA; /* :1507-1509 */
exit:
  res->new_tail = tail; /* :1520 */
  res->n_records = n;
  res->n_rows = n_rows;
  res->ret = err >= 0 ? (int64_t)n : err; /* :1521 */
  res->carry_tstamp = carry ? carry->tstamp : INVALID_TS;
  if (sw) res->num_unknown_count_interval = carry->last[0];
  return 0;
}

int hbt_oracle_consume(const uint8_t* ring, uint64_t data_size, uint64_t tail, uint64_t head,
                       const hbt_decode_desc* d, hbt_oracle_carry* carry, uint64_t* rows,
                       uint64_t rows_cap, hbt_oracle_hist* hist, hbt_cpu_result* res) {
  return consume_core(ring, data_size, tail, head, d, carry, rows, rows_cap, hist, res, 0);
}

/* Same with the CPU id that HBT_AGG_SWITCH_EVENTS writes into Event::comp_unit_id
 * (comp_unit_id_ = cpu, PerCpuThreadSwitchGenerator.h:199).  rows then holds 16-byte
 * hbt_switch_event records, rows_cap counts events; carry->tstamp is last_tstamp_
 * (0 at start: numeric_limits<TimeStamp>::min(), :437) and carry->last[0] in_write_errors_. */
int hbt_oracle_consume_cpu(const uint8_t* ring, uint64_t data_size, uint64_t tail, uint64_t head,
                           const hbt_decode_desc* d, hbt_oracle_carry* carry, uint64_t* rows,
                           uint64_t rows_cap, hbt_oracle_hist* hist, hbt_cpu_result* res,
                           uint32_t cpu) {
  return consume_core(ring, data_size, tail, head, d, carry, rows, rows_cap, hist, res, cpu);
}

/* ---- counter groups ------------------------------------------------------ */

/* GroupReadValues::getCount, CpuEventsGroup.h:449-459. */
static uint64_t get_count(uint64_t count, uint64_t time_enabled, uint64_t time_running) {
  if (time_enabled == 0) return 0;
  return (uint64_t)((double)count * (double)time_enabled / (double)time_running);
}

/* PerCpuBase<T>::read (PerCpuBase.h:104-119: zero, then read+accum per CPU;
 * a failed CPU read fails the whole read), GroupReadValues::zero/accum
 * (CpuEventsGroup.h:508-522), getCount (:449-459), PerfMonitor::log's rate
 * expression (dynolog/src/PerfMonitor.cpp:52-67), ratio convention
 * double(getCount(a))/double(getCount(b)), MetricSeries::diff
 * (dynolog/src/metric_frame/MetricSeries.h:221-227) between consecutive
 * intervals.  Layouts as hbt_counts_pipeline_dev.  Outputs may be NULL. */
int hbt_oracle_counts(const uint64_t* reads, uint64_t n_intervals, uint32_t n_cpus,
                      uint32_t n_events, const hbt_ratio* ratios, uint32_t n_ratios,
                      uint64_t* sums, uint64_t* scaled, float* rates, double* ratio_out,
                      uint64_t* deltas) {
  if (n_events == 0 || n_events > HBT_MAX_EVENTS) return -EINVAL;
  const uint64_t W = 3 + (uint64_t)n_events;
  uint64_t acc[3 + HBT_MAX_EVENTS];
  uint64_t cur[HBT_MAX_EVENTS], prev[HBT_MAX_EVENTS];
  memset(prev, 0, sizeof(prev));
  for (uint64_t i = 0; i < n_intervals; ++i) {
    int ok = 1;
    memset(acc, 0, sizeof(acc)); /* zero(), :508-512 */
    acc[0] = n_events;
    for (uint32_t cpu = 0; cpu < n_cpus; ++cpu) {
      const uint64_t* t = reads + (i * n_cpus + cpu) * W;
      if (t[0] != n_events) ok = 0; /* failed read, CpuEventsGroup.h:599-614 */
      for (uint64_t w = 1; w < W; ++w) acc[w] += t[w]; /* accum(), :514-522 */
    }
    if (!ok) acc[0] = 0;
    if (sums) memcpy(sums + i * W, acc, 8 * W);
    for (uint32_t e = 0; e < n_events; ++e) {
      uint64_t count = ok ? get_count(acc[3 + e], acc[1], acc[2]) : 0;
      cur[e] = count;
      if (scaled) scaled[i * n_events + e] = count;
      if (rates) /* PerfMonitor.cpp:56-57 */
        rates[i * n_events + e] = (float)((double)count * 1e3 / (double)acc[2]);
      if (deltas) deltas[i * n_events + e] = count - prev[e];
    }
    if (ratio_out)
      for (uint32_t r = 0; r < n_ratios; ++r)
        ratio_out[i * n_ratios + r] =
            (double)cur[ratios[r].num_event] / (double)cur[ratios[r].den_event];
    memcpy(prev, cur, sizeof(prev));
  }
  return 0;
}
