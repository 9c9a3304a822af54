// k_walk: the decode for streams of MIXED record types and sizes (context-switch rings, config 4's
// mixed stream), where the uniform fast path of k_decode has nothing to hold on to.
//
// Replaces the same reference code as k_decode (CpuEventsGroup::consume, CpuEventsGroup.h:1326-1522,
// and the generators' handlers: PerCpuCountSampleGenerator.h:122-195, PerCpuThreadSwitchGenerator.h:
// 228-412) and produces the same SegResult / row slabs, so k_finalize and everything behind it are
// shared.
//
// Why a second kernel.  k_decode's general path gives every THREAD a 64-byte sub-chunk of a staged
// tile and lets it scan all of its words for a plausible header: the work is proportional to the
// bytes (~28 issue cycles per record on config 4's streams).  Here one WARP owns a segment and works
// through it in rounds of 8 KiB, staged in the warp's own slice of shared memory; each lane owns a
// 256-byte chunk of the round, guesses the first record start in it (first header whose chain of
// successors stays plausible) and then simply follows header.size from record to record -- work
// proportional to the RECORDS.  The seams between the 32 lanes are settled inside the warp: lane l's
// walk must begin where the records before it end; a lane whose guess was wrong walks again from the
// right place.  What leaves a round is exact, and the position after its last record carries into
// the next round, so only a segment's very first record remains a guess that k_finalize has to
// confirm (as for k_decode).  No block barriers: warps of a CTA share nothing.
//
// Round = stage, pass A (guess + walk, counting record types on the way), seam resolution, pass B
// (a second walk over the now exact records for the outputs that need stream-order indices:
// proto-events, delta rows).
//
// Shared-memory layout: chunk j of the round lives at j * (256 + 8) bytes, so that lanes reading the
// same offset of their own chunks (the common case: every lane starts at its chunk's first header)
// fall on different banks.  One extra chunk of lookahead is staged for the records of the last
// lane; reads beyond it (a long record's tail) go to global memory.
#pragma once
#include "decode.cuh"

namespace hbt {

constexpr int kWalkChunk = 256;                 // bytes a lane owns per round
constexpr int kWalkRound = 32 * kWalkChunk;     // bytes a warp covers per round
constexpr int kWalkPitch = kWalkChunk + 8;      // distance of two chunks in shared memory
constexpr int kWalkStaged = 33;                 // chunks staged per round (the round + lookahead)
constexpr int kWalkWarps = 4;                   // warps (= segments) per CTA
constexpr int kWalkSmemPerWarp = kWalkStaged * kWalkPitch;
static_assert(kWalkChunk == 256, "WalkView::word shifts by 8");

struct WalkView {
  uint32_t s;      // shared-space address of the warp's staging slice
  uint64_t base;   // ring position of its first byte
  const RingDesc* ring;
  __device__ __forceinline__ uint64_t word(uint64_t pos) const {
    const uint64_t rel = pos - base;  // (positions before the round wrap to huge values)
    if (rel < (uint64_t)(kWalkStaged * kWalkChunk)) {
      const uint32_t r = (uint32_t)rel;
      return lds_u64(s + r + ((r >> 8) << 3));
    }
    return ring_word_g(*ring, pos);
  }
};

struct WalkLane {
  uint64_t entry;    // first record start in my chunk (kNone: none)
  uint64_t exit;     // where my walk left the chunk
  uint64_t err_pos;
  int32_t err;       // error met by the walk (only meaningful once the entry is confirmed)
  uint32_t n_rec;
  uint32_t n_handled;      // SWITCH_EVENTS: records that emit a proto-event
  uint32_t n_lost;         // handled LOST records
  uint64_t lost_sum;       // their num_lost
  uint64_t cnt_lo, cnt_hi; // per-type counts, 8 bits each: counter index 0..7 / 8..10
  // COUNT_DELTAS: the SAMPLE / LOST events of my chunk
  uint64_t first_ev, last_ev;  // encoded (ev_make), 0 = none
  uint64_t last_sample;        // kNone = none
  uint32_t n_rows;             // SAMPLEs whose previous event (in my chunk) is a SAMPLE
  uint32_t n_unknown;          // SAMPLEs whose previous event (in my chunk) is a LOST
};

__device__ __forceinline__ void walk_clear(WalkLane& w) {
  w.err = 0;
  w.err_pos = 0;
  w.n_rec = w.n_handled = w.n_lost = 0;
  w.lost_sum = 0;
  w.cnt_lo = w.cnt_hi = 0;
  w.first_ev = w.last_ev = 0;
  w.last_sample = kNone;
  w.n_rows = w.n_unknown = 0;
}

// Follow header.size from `from` while records start before `end`, counting per type.  The walk
// must be harmless on garbage (a guess may be wrong): it stops at the first header that cannot be
// one, or that the decode mode refuses, and remembers why.
template <int AGG>
__device__ __forceinline__ void walk_chunk(const hbt_decode_desc& d, const WalkView& v, uint64_t from, uint64_t end,
                                           uint32_t tmask, WalkLane& w) {
  walk_clear(w);
  uint32_t k = EV_NONE;  // kind of the last SAMPLE / LOST event of this walk
  uint64_t pos = from;
  while (pos < end) {
    const uint64_t h = v.word(pos);
    const uint32_t type = hdr_type(h), size = hdr_size(h);
    if (size < 8 || (size & 7)) {
      w.err = -74;  // -EBADMSG
      w.err_pos = pos;
      break;
    }
    const int ci = counter_of(type, tmask);
    if (ci >= 0) {
      if (type == REC_SAMPLE || type == REC_READ) {
        // hard size checks of consume() (CpuEventsGroup.h:1435-1440, 1451-1456)
        const bool group = type == REC_READ || d.sample_layout == HBT_SAMPLE_GROUP;
        const uint64_t want = group ? 40 + 8 * v.word(pos + 16) : d.ip_sample_bytes;
        if ((uint64_t)size != want) {
          w.err = -1;  // -EPERM
          w.err_pos = pos;
          break;
        }
      }
      if (AGG == HBT_AGG_SWITCH_EVENTS) {
        if (size < sw_min_size(type)) {  // the handler would read past the record
          w.err = -1;
          w.err_pos = pos;
          break;
        }
        ++w.n_handled;
      }
      if (AGG == HBT_AGG_COUNT_DELTAS && type == REC_SAMPLE && d.sample_layout == HBT_SAMPLE_GROUP) {
        if (k == EV_NONE) w.first_ev = ev_make(pos, EV_SAMPLE);
        else if (k == EV_LOST) ++w.n_unknown;
        else ++w.n_rows;
        k = EV_SAMPLE;
        w.last_ev = ev_make(pos, EV_SAMPLE);
        w.last_sample = pos;
      }
      if (ci == HBT_CNT_LOST) {
        w.lost_sum += v.word(pos + 16);
        ++w.n_lost;
        if (AGG == HBT_AGG_COUNT_DELTAS) {
          if (k == EV_NONE) w.first_ev = ev_make(pos, EV_LOST);
          k = EV_LOST;
          w.last_ev = ev_make(pos, EV_LOST);
        }
      } else if (ci < 8) {
        w.cnt_lo += 1ull << (8 * ci);
      } else {
        w.cnt_hi += 1ull << (8 * (ci - 8));
      }
    }
    ++w.n_rec;
    pos += size;
  }
  w.exit = pos;
}

// First record start at or after `from` (below `end`) judged by looks: a plausible header whose next
// three successors are plausible too (or lie beyond the window).  kNone if there is none.
__device__ __forceinline__ uint64_t guess_entry(const WalkView& v, uint64_t from, uint64_t end, uint64_t head) {
  for (uint64_t q = from; q < end; q += 8) {
    uint64_t h = v.word(q);
    if (!plausible(h)) continue;
    uint64_t nx = q + hdr_size(h);
    bool ok = true;
#pragma unroll 1
    for (int k = 0; k < 3 && nx < head; ++k) {
      h = v.word(nx);
      if (!plausible(h)) {
        ok = false;
        break;
      }
      nx += hdr_size(h);
    }
    if (ok) return q;
  }
  return kNone;
}

__device__ __forceinline__ uint64_t warp_max_scan_incl(uint64_t x, uint32_t lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint64_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= (unsigned)o) x = max(x, y);
  }
  return x;
}
__device__ __forceinline__ uint32_t warp_sum_scan_incl(uint32_t x, uint32_t lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= (unsigned)o) x += y;
  }
  return x;
}
__device__ __forceinline__ uint64_t warp_sum64(uint64_t x) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
  return x;
}
__device__ __forceinline__ uint64_t warp_max64(uint64_t x) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) x = max(x, __shfl_xor_sync(0xffffffffu, x, o));
  return x;
}
__device__ __forceinline__ void sts_u64(uint32_t saddr, uint64_t v) {
  asm volatile("st.shared.u64 [%0], %1;" ::"r"(saddr), "l"(v) : "memory");
}

template <int AGG>
__global__ void __launch_bounds__(32 * kWalkWarps, 6) k_walk(const DecodeParams p, uint32_t n_launch) {
  extern __shared__ __align__(16) uint8_t walk_smem[];
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t gw = blockIdx.x * kWalkWarps + warp;
  if (gw >= n_launch) return;
  const uint32_t seg_index = p.run_list ? p.run_list[gw] : gw;
  const SegDesc seg = p.segs[seg_index];
  const RingDesc ring = p.rings[seg.cpu];
  const hbt_decode_desc& d = p.d;
  const uint32_t tmask = type_mask(d.handled_mask);
  constexpr unsigned FULL = 0xffffffffu;
  const bool ev_samples = AGG == HBT_AGG_COUNT_DELTAS && d.sample_layout == HBT_SAMPLE_GROUP;
  const uint32_t row_words = 2 + d.n_events;

  WalkView v;
  v.s = smem_u32(walk_smem) + warp * kWalkSmemPerWarp;
  v.ring = &ring;

  // records are decoded while they START in [seg_lo, seg_limit)
  const uint64_t seg_lo = max(seg.tile_begin * (uint64_t)kTile, ring.tail);
  const uint64_t seg_limit = min(seg.tile_end * (uint64_t)kTile, ring.head);

  // exact next record start, or kNone while the segment still guesses its entry
  uint64_t pos = seg.seg == 0 ? ring.tail : seg.forced_entry;
  uint64_t entry = kNone;  // first record start decoded (k_finalize checks it against the chain)
  int32_t err = 0;
  uint64_t err_pos = 0;
  // warp-uniform running state
  uint64_t n_rows = 0;           // rows / proto-events written so far
  uint64_t ev = 0;               // last SAMPLE / LOST event so far (encoded)
  uint64_t last_sample = kNone;
  uint64_t seg_first_ev = 0;     // the segment's first SAMPLE when no event precedes it (the seam)
  // per-lane accumulators, reduced when the segment is done
  uint32_t acc_rec = 0, acc_lost_n = 0, acc_unknown = 0;
  uint64_t acc_lost = 0;
  uint64_t acc_lo = 0, acc_hi = 0;  // packed 8-bit counts of the rounds since the last flush
  uint32_t acc_rounds = 0;
  uint64_t total = 0;               // lane c holds the segment's count of counter c
  auto flush = [&]() {
#pragma unroll
    for (int c = 0; c < HBT_NUM_COUNTERS; ++c) {
      if (c == HBT_CNT_LOST) continue;
      const uint32_t x = c < 8 ? (uint32_t)(acc_lo >> (8 * c)) & 0xFFu : (uint32_t)(acc_hi >> (8 * (c - 8))) & 0xFFu;
      const uint32_t s = __reduce_add_sync(FULL, x);
      if (lane == (unsigned)c) total += s;
    }
    acc_lo = acc_hi = 0;
    acc_rounds = 0;
  };

  for (uint64_t base = seg_lo & ~(uint64_t)(kWalkRound - 1); base < seg_limit && err == 0; base += kWalkRound) {
    // ---- stage the round (+ lookahead): line j of the copy is chunk j -----------------------------
    __syncwarp();
    v.base = base;
#pragma unroll 11
    for (int j = 0; j < kWalkStaged; ++j) {
      const uint64_t x = ld_stream_u64(
          reinterpret_cast<const uint64_t*>(ring.base + ((base + (uint64_t)j * kWalkChunk + lane * 8) & ring.mask)));
      sts_u64(v.s + j * kWalkPitch + lane * 8, x);
    }
    __syncwarp();

    // my chunk, clipped to the segment
    const uint64_t cs = max(base + (uint64_t)lane * kWalkChunk, seg_lo);
    const uint64_t ce = min(base + (uint64_t)(lane + 1) * kWalkChunk, seg_limit);
    const bool have = cs < ce;

    // ---- pass A: guess and walk -------------------------------------------------------------
    WalkLane w;
    walk_clear(w);
    w.entry = kNone;
    w.exit = kNone;
    if (have) {
      if (pos != kNone && pos >= cs && pos < ce) w.entry = pos;   // the exact entry lies in my chunk
      else if (pos != kNone && pos >= ce) w.entry = kNone;        // a record spans my whole chunk
      else w.entry = guess_entry(v, cs, ce, ring.head);           // (the host's entry_hint is of no use here: it
                                                                  //  may name a record that is not the FIRST one)
      if (w.entry != kNone) walk_chunk<AGG>(d, v, w.entry, ce, tmask, w);
    }

    // ---- seams: lane l must start where the records before it end --------------------------------
    uint64_t cur;
    {
      // if every guess is right: the chain position in front of my chunk is the exit of the nearest
      // walk before me (exits grow with the lane), else what the round inherited
      const uint32_t x = w.entry != kNone ? (uint32_t)(w.exit - base) + 1u : 0u;
      uint32_t incl = x;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(FULL, incl, o);
        if (lane >= (unsigned)o) incl = max(incl, y);
      }
      uint32_t excl = __shfl_up_sync(FULL, incl, 1);
      if (lane == 0) excl = 0;
      const uint64_t before = excl ? base + excl - 1 : pos;
      bool ok;
      if (before == kNone) ok = true;                       // nothing known in front of me
      else if (!have || before >= ce) ok = w.entry == kNone;
      else ok = w.entry == before;
      ok = ok && w.err == 0;
      if (__all_sync(FULL, ok)) {
        const uint32_t last = __shfl_sync(FULL, incl, 31);
        cur = last ? base + last - 1 : pos;
      } else {
        // some guess was wrong (or a walk met an error): settle the lanes one after the other
        cur = pos;
        for (int l = 0; l < 32; ++l) {
          const uint64_t l_cs = max(base + (uint64_t)l * kWalkChunk, seg_lo);
          const uint64_t l_ce = min(base + (uint64_t)(l + 1) * kWalkChunk, seg_limit);
          if (l_cs >= l_ce) continue;
          if (cur == kNone) {
            // the segment's entry is still a guess: the first lane that found a start supplies it
            const uint64_t e = __shfl_sync(FULL, w.entry, l);
            if (e == kNone) continue;
            cur = e;
          }
          if (cur >= l_ce) {
            // no record starts in lane l's chunk: whatever it guessed was not a record
            if (lane == l) w.entry = kNone;
            continue;
          }
          if (lane == l && w.entry != cur) {  // wrong (or no) guess: walk again from the right place
            w.entry = cur;
            walk_chunk<AGG>(d, v, cur, ce, tmask, w);
          }
          const int32_t e_l = __shfl_sync(FULL, w.err, l);
          cur = __shfl_sync(FULL, w.exit, l);
          if (e_l != 0) {  // a real error on the confirmed chain: consume() stops in front of that record
            err = e_l;
            err_pos = __shfl_sync(FULL, w.err_pos, l);
            if ((int)lane > l) w.entry = kNone;  // records consume() never reaches
            break;
          }
        }
      }
    }
    if (entry == kNone) {
      uint64_t m = w.entry;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) m = min(m, __shfl_xor_sync(FULL, m, o));
      entry = m;
    }
    if (w.entry == kNone) walk_clear(w);
    const uint64_t stop = w.err ? w.err_pos : ce;  // my records start in [w.entry, stop)

    // ---- pass B: outputs that need stream-order indices ------------------------------------------
    if (AGG == HBT_AGG_SWITCH_EVENTS) {
      const uint32_t incl = warp_sum_scan_incl(w.n_handled, lane);
      uint64_t row_i = n_rows + (incl - w.n_handled);
      if (w.n_handled) {
        uint64_t q = w.entry;
        while (q < stop) {
          const uint64_t h = v.word(q);
          if (counted(hdr_type(h), tmask)) {
            const ulonglong2 e =
                sw_proto(h, d.tid_level, p.cpu_base + seg.cpu, [&](uint32_t off) { return v.word(q + off); });
            if (row_i < seg.rows_cap) *reinterpret_cast<ulonglong2*>(seg.rows + (1 + row_i) * 2) = e;
            ++row_i;
          }
          q += hdr_size(h);
        }
      }
      n_rows += __shfl_sync(FULL, incl, 31);
    }
    if (AGG == HBT_AGG_COUNT_DELTAS) {
      // the event in front of my chunk: last event of the nearest lane before me that has one, else
      // what the segment has seen so far
      const uint64_t incl_ev = warp_max_scan_incl(w.last_ev, lane);
      uint64_t in_ev = __shfl_up_sync(FULL, incl_ev, 1);
      if (lane == 0 || in_ev == 0) in_ev = ev;
      uint32_t my_rows, my_unknown;
      uint64_t my_seam;
      auto derive = [&]() {
        my_rows = w.n_rows;
        my_unknown = w.n_unknown;
        my_seam = 0;
        if (ev_kind(w.first_ev) == EV_SAMPLE) {
          const uint32_t k = ev_kind(in_ev);
          if (k == EV_SAMPLE) ++my_rows;                // interval from the previous SAMPLE
          else if (k == EV_LOST) ++my_unknown;          // re-seed, PerCpuCountSampleGenerator.h:135-154
          else my_seam = w.first_ev;                    // nothing before it in this segment: k_finalize's seam
        }
      };
      derive();
      const uint32_t incl_r = warp_sum_scan_incl(my_rows, lane);
      const uint64_t row0 = n_rows + (incl_r - my_rows);
      int32_t late_err = 0;
      uint64_t late_pos = 0;
      if (w.last_ev != 0) {
        bool valid = ev_kind(in_ev) == EV_SAMPLE;
        uint64_t ppos = ev_pos(in_ev);
        uint64_t row_i = row0;
        uint64_t q = w.entry;
        while (q < stop) {
          const uint64_t h = v.word(q);
          const uint32_t type = hdr_type(h);
          if (counted(type, tmask)) {
            if (type == REC_SAMPLE && ev_samples) {
              if (valid) {
                const uint64_t ts = v.word(q + 8), pts = v.word(ppos + 8);
                if (pts > ts) {  // :156-163: the interval is refused, consume() stops in front of the record
                  late_err = -1;
                  late_pos = q;
                  break;
                }
                if (row_i >= seg.rows_cap) {  // (only with nr != n_events records) a full output ring, :191-194
                  late_err = -28;
                  late_pos = q;
                  break;
                }
                uint64_t* row = seg.rows + (1 + row_i) * row_words;
                row[0] = pts;
                row[1] = ts - pts;
                for (uint32_t i = 0; i < d.n_events; ++i)
                  row[2 + i] = xform(d, i, v.word(q + 40 + 8ull * i)) - xform(d, i, v.word(ppos + 40 + 8ull * i));
                ++row_i;
              }
              valid = true;
              ppos = q;
            } else if (type == REC_LOST) {
              valid = false;
            }
          }
          q += hdr_size(h);
        }
      }
      const uint32_t bad = __ballot_sync(FULL, late_err != 0);
      uint32_t total_rows = __shfl_sync(FULL, incl_r, 31);
      if (bad) {
        // consume() stops in front of that SAMPLE: what my walk counted behind it, and every lane
        // behind me, is dropped
        const int l = __ffs(bad) - 1;
        err = __shfl_sync(FULL, late_err, l);
        err_pos = __shfl_sync(FULL, late_pos, l);
        if ((int)lane == l) {
          walk_chunk<AGG>(d, v, w.entry, late_pos, tmask, w);
          derive();
          w.last_ev = ev_make(late_pos, EV_LOST);  // count_.tstamp = invalid
        } else if ((int)lane > l) {
          w.entry = kNone;
          walk_clear(w);
          my_rows = my_unknown = 0;
          my_seam = 0;
        }
        total_rows = __reduce_add_sync(FULL, my_rows);
        cur = err_pos;
      }
      n_rows += total_rows;
      acc_unknown += my_unknown;
      const uint64_t le = warp_max64(w.last_ev);
      if (le) ev = le;
      const uint64_t ls = warp_max64(w.last_sample == kNone ? 0 : w.last_sample + 1);
      if (ls) last_sample = ls - 1;
      if (seg_first_ev == 0) seg_first_ev = warp_max64(my_seam);
    }

    // ---- keep the lanes' counts ------------------------------------------------------------------
    acc_rec += w.n_rec;
    acc_lost_n += w.n_lost;
    acc_lost += w.lost_sum;
    acc_lo += w.cnt_lo;
    acc_hi += w.cnt_hi;
    if (++acc_rounds == 7) flush();  // 7 rounds x 32 records per lane stay below 256
    pos = cur;  // exact from here on (kNone only if nothing at all was found yet)
  }

  flush();
  const uint64_t n_rec = warp_sum64(acc_rec);
  const uint32_t n_lost = __reduce_add_sync(FULL, acc_lost_n);
  const uint64_t lost_sum = warp_sum64(acc_lost);
  const uint32_t n_unknown = __reduce_add_sync(FULL, acc_unknown);
  SegResult& r = p.results[seg_index];
  if (lane < HBT_NUM_COUNTERS) r.counters[lane] = lane == HBT_CNT_LOST ? lost_sum : total;
  if (lane == 0) {
    r.entry = entry;
    r.exit = pos;
    r.n_records = n_rec;
    r.n_rows = n_rows;
    r.n_unknown = n_unknown;
    r.first_ev = seg_first_ev;
    r.last_ev = ev;
    r.last_sample = last_sample;
    r.err = err;
    r.err_pos = err_pos;
    r.slow_tiles = 0;
    r.hist_overflow = 0;
    r.n_lost_recs = n_lost;
  }
}

}  // namespace hbt
