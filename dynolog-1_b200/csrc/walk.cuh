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
// fall on different banks.  Two extra chunks of lookahead are staged for the records of the last
// lane; reads beyond them (a long record's tail) go to global memory.  Inside a round every position
// is a 32-bit offset from the round's first byte.
#pragma once
#include "decode.cuh"

namespace hbt {

constexpr int kWalkChunk = 256;                 // bytes a lane owns per round
constexpr int kWalkRound = 32 * kWalkChunk;     // bytes a warp covers per round
constexpr int kWalkPitch = kWalkChunk + 8;      // distance of two chunks in shared memory
constexpr int kWalkStaged = 34;                 // chunks staged per round (the round + lookahead)
constexpr uint32_t kWalkStagedBytes = kWalkStaged * kWalkChunk;
constexpr int kWalkWarps = 4;                   // warps (= segments) per CTA
constexpr int kWalkStageSmem = kWalkStaged * kWalkPitch;
// COUNT_DELTAS: {current, previous} SAMPLE of every row of the round, in row order (a SAMPLE is at
// least 40 bytes even with nr = 0)
constexpr int kWalkMaxRows = kWalkRound / 40 + 3;
constexpr uint32_t kFarPrev = 0xFFFFFFFEu;      // "previous SAMPLE lies before this round"
static_assert(kWalkChunk == 256, "WalkView shifts by 8");
static_assert(kWalkStageSmem % 8 == 0, "row descriptors follow the staging slice");

__host__ __device__ constexpr int walk_smem_per_warp(int agg) {
  return kWalkStageSmem + (agg == HBT_AGG_COUNT_DELTAS ? kWalkMaxRows * 8 : 0);
}

struct WalkView {
  uint32_t s;      // shared-space address of the warp's staging slice
  uint64_t base;   // ring position of its first byte
  const RingDesc* ring;
  // word at offset r of the round, r inside the staged bytes (every header of the round, and fields
  // up to 512 bytes behind one)
  __device__ __forceinline__ uint64_t near(uint32_t r) const { return lds_u64(s + r + ((r >> 8) << 3)); }
  // any offset
  __device__ __forceinline__ uint64_t far(uint32_t r) const {
    return r < kWalkStagedBytes ? near(r) : ring_word_g(*ring, base + r);
  }
  // any ring position
  __device__ __forceinline__ uint64_t at(uint64_t pos) const {
    const uint64_t rel = pos - base;  // (positions before the round wrap to huge values)
    return rel < (uint64_t)kWalkStagedBytes ? near((uint32_t)rel) : ring_word_g(*ring, pos);
  }
};

// events of a chunk, round-relative: (offset + 1) << 2 | kind, 0 = none
__device__ __forceinline__ uint32_t rev_make(uint32_t r, uint32_t kind) { return ((r + 1) << 2) | kind; }
__device__ __forceinline__ uint32_t rev_off(uint32_t e) { return (e >> 2) - 1; }

struct WalkLane {
  uint32_t entry;    // first record start in my chunk (kNoRel: none)
  uint32_t exit;     // where my walk left the chunk
  uint32_t err_r;
  int32_t err;       // error met by the walk (only meaningful once the entry is confirmed)
  uint32_t n_rec;
  uint32_t n_handled;      // SWITCH_EVENTS: records that emit a proto-event
  uint32_t n_lost;         // handled LOST records
  uint64_t lost_sum;       // their num_lost
  uint64_t cnt_lo, cnt_hi; // per-type counts, 8 bits each: counter index 0..7 / 8..10
  // COUNT_DELTAS: the SAMPLE / LOST events of my chunk
  uint32_t first_ev, last_ev;  // rev_make, 0 = none
  uint32_t last_sample;        // offset + 1, 0 = none
  uint32_t n_rows;             // SAMPLEs whose previous event (in my chunk) is a SAMPLE
  uint32_t n_unknown;          // SAMPLEs whose previous event (in my chunk) is a LOST
};

__device__ __forceinline__ void walk_clear(WalkLane& w) {
  w.err = 0;
  w.err_r = 0;
  w.n_rec = w.n_handled = w.n_lost = 0;
  w.lost_sum = 0;
  w.cnt_lo = w.cnt_hi = 0;
  w.first_ev = w.last_ev = 0;
  w.last_sample = 0;
  w.n_rows = w.n_unknown = 0;
}

// Follow header.size from offset `from` while records start before `end`, counting per type.  The
// walk must be harmless on garbage (a guess may be wrong): it stops at the first header that cannot
// be one, or that the decode mode refuses, and remembers why.
template <int AGG>
__device__ __forceinline__ void walk_chunk(const hbt_decode_desc& d, const WalkView& v, uint32_t from, uint32_t end,
                                           uint32_t tmask, WalkLane& w) {
  walk_clear(w);
  const bool group_layout = d.sample_layout == HBT_SAMPLE_GROUP;
  uint32_t k = EV_NONE;  // kind of the last SAMPLE / LOST event of this walk
  uint32_t r = from;
  while (r < end) {
    const uint64_t h = v.near(r);
    const uint32_t type = hdr_type(h), size = hdr_size(h);
    if (size < 8 || (size & 7)) {
      w.err = -74;  // -EBADMSG
      w.err_r = r;
      break;
    }
    if (counted(type, tmask)) {
      const uint32_t ci = (uint32_t)(kCiNib >> (4 * type)) & 15u;
      if (type == REC_SAMPLE || type == REC_READ) {
        // hard size checks of consume() (CpuEventsGroup.h:1435-1440, 1451-1456)
        const bool group = type == REC_READ || group_layout;
        const uint64_t want = group ? 40 + 8 * v.near(r + 16) : d.ip_sample_bytes;
        if ((uint64_t)size != want) {
          w.err = -1;  // -EPERM
          w.err_r = r;
          break;
        }
      }
      if (AGG == HBT_AGG_SWITCH_EVENTS) {
        if (size < sw_min_size(type)) {  // the handler would read past the record
          w.err = -1;
          w.err_r = r;
          break;
        }
        ++w.n_handled;
      }
      if (AGG == HBT_AGG_COUNT_DELTAS && type == REC_SAMPLE && group_layout) {
        if (k == EV_NONE) w.first_ev = rev_make(r, EV_SAMPLE);
        else if (k == EV_LOST) ++w.n_unknown;
        else ++w.n_rows;
        k = EV_SAMPLE;
        w.last_ev = rev_make(r, EV_SAMPLE);
        w.last_sample = r + 1;
      }
      if (type == REC_LOST) {
        w.lost_sum += v.near(r + 16);
        ++w.n_lost;
        if (AGG == HBT_AGG_COUNT_DELTAS) {
          if (k == EV_NONE) w.first_ev = rev_make(r, EV_LOST);
          k = EV_LOST;
          w.last_ev = rev_make(r, EV_LOST);
        }
      } else if (ci < 8) {
        w.cnt_lo += 1ull << (8 * ci);
      } else {
        w.cnt_hi += 1ull << (8 * (ci - 8));
      }
    }
    ++w.n_rec;
    r += size;
  }
  w.exit = r;
}

// First record start at or after offset `from` (below `end`) judged by looks: a plausible header
// whose next three successors are plausible too (or lie beyond the window).  kNoRel if there is none.
__device__ __forceinline__ uint32_t guess_entry(const WalkView& v, uint32_t from, uint32_t end, uint32_t head_r) {
  for (uint32_t q = from; q < end; q += 8) {
    uint64_t h = v.near(q);
    if (!plausible(h)) continue;
    uint32_t nx = q + hdr_size(h);
    bool ok = true;
#pragma unroll 1
    for (int k = 0; k < 3 && nx < head_r; ++k) {
      h = v.far(nx);
      if (!plausible(h)) {
        ok = false;
        break;
      }
      nx += hdr_size(h);
    }
    if (ok) return q;
  }
  return kNoRel;
}

__device__ __forceinline__ uint32_t warp_max_scan_incl(uint32_t x, uint32_t lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
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
__device__ __forceinline__ void sts_u64(uint32_t saddr, uint64_t v) {
  asm volatile("st.shared.u64 [%0], %1;" ::"r"(saddr), "l"(v) : "memory");
}
__device__ __forceinline__ void sts_u32x2(uint32_t saddr, uint32_t a, uint32_t b) {
  asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(saddr), "r"(a), "r"(b) : "memory");
}
__device__ __forceinline__ uint2 lds_u32x2(uint32_t saddr) {
  uint2 r;
  asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(r.x), "=r"(r.y) : "r"(saddr));
  return r;
}
__device__ __forceinline__ ulonglong2 ldg_stream_u64x2(const void* p) {
  ulonglong2 r;
  asm volatile("ld.global.cs.v2.u64 {%0, %1}, [%2];" : "=l"(r.x), "=l"(r.y) : "l"(p));
  return r;
}

// The proto-event of a handled ContextSwitch record, without branches on the record type (the lanes
// of a warp look at records of different types): same result as sw_proto.
__device__ __forceinline__ ulonglong2 sw_proto_flat(const WalkView& v, uint32_t q, uint64_t h, uint32_t level,
                                                     uint32_t cu) {
  const uint32_t type = hdr_type(h), size = hdr_size(h), misc = (uint32_t)(h >> 32) & 0xFFFF;
  const bool lost = type == REC_LOST, comm = type == REC_COMM, ex = type == REC_EXIT, fork = type == REC_FORK;
  const bool sw = !(lost || comm || ex || fork);
  const bool out = sw && (misc & kMiscSwitchOut);
  // the record's own time: LOST +32 (sample_id), COMM last word, the others +24
  const uint64_t ts = v.far(q + (lost ? 32u : comm ? size - 8u : 24u));
  const uint64_t a = v.near(q + (comm ? 8u : 16u));
  const uint32_t tag = (lost || out) ? (uint32_t)HBT_TAG_NA : (ex || fork) ? (uint32_t)a : (uint32_t)(a >> 32);
  const uint32_t et = lost   ? HBT_SWEV_WRITE_ERRORS_START
                      : comm ? HBT_SWEV_COMM
                      : ex   ? HBT_SWEV_THREAD_DESTRUCTION
                      : fork ? HBT_SWEV_THREAD_CREATION
                      : !out ? HBT_SWEV_SWITCH_IN
                      : (misc & kMiscSwitchOutPreempt) ? HBT_SWEV_SWITCH_OUT_PREEMPT
                                                       : HBT_SWEV_SWITCH_OUT_YIELD;
  return make_ulonglong2(ts, sw_word1(et, lost ? (uint32_t)HBT_LEVEL_INVALID : level, cu, tag));
}

template <int AGG>
__global__ void __launch_bounds__(32 * kWalkWarps, AGG == HBT_AGG_COUNT_DELTAS ? 5 : 6)
    k_walk(const DecodeParams p, uint32_t n_launch) {
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
  // cooperative row writer: lane -> (row of the batch, word of the row)
  const uint32_t rows_per_iter = row_words <= 32 ? 32 / row_words : 0;
  const uint32_t sub_row = rows_per_iter ? lane / row_words : 0;
  const uint32_t sub_col = rows_per_iter ? lane - sub_row * row_words : lane;

  WalkView v;
  v.s = smem_u32(walk_smem) + warp * walk_smem_per_warp(AGG);
  v.ring = &ring;
  const uint32_t s_desc = v.s + kWalkStageSmem;  // COUNT_DELTAS: row descriptors
  const bool src16 = (reinterpret_cast<uintptr_t>(ring.base) & 15) == 0;

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
  uint64_t ev = 0;               // last SAMPLE / LOST event so far (ev_make)
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
    // ---- stage the round (+ lookahead) ------------------------------------------------------------
    __syncwarp();
    v.base = base;
    if (src16) {
      // one 16-byte load per lane covers two chunks per instruction
      const uint32_t c0 = lane >> 4, off = (lane & 15) * 16;
#pragma unroll
      for (int j = 0; j < kWalkStaged / 2; ++j) {
        const uint32_t c = 2 * j + c0;
        const ulonglong2 x = ldg_stream_u64x2(ring.base + ((base + c * kWalkChunk + off) & ring.mask));
        sts_u64(v.s + c * kWalkPitch + off, x.x);
        sts_u64(v.s + c * kWalkPitch + off + 8, x.y);
      }
    } else {
#pragma unroll 17
      for (int j = 0; j < kWalkStaged; ++j) {
        const uint64_t x = ld_stream_u64(
            reinterpret_cast<const uint64_t*>(ring.base + ((base + (uint64_t)j * kWalkChunk + lane * 8) & ring.mask)));
        sts_u64(v.s + j * kWalkPitch + lane * 8, x);
      }
    }
    __syncwarp();

    // round-relative geometry: the segment's part of the round, my chunk's part of that
    const uint32_t lo_r = seg_lo > base ? (uint32_t)(seg_lo - base) : 0u;
    const uint32_t lim_r = (uint32_t)min(seg_limit - base, (uint64_t)kWalkRound);
    const uint32_t head_r = (uint32_t)min(ring.head - base, (uint64_t)0xFFFF0000u);
    const uint32_t cs = max(lane * kWalkChunk, lo_r);
    const uint32_t ce = min((lane + 1) * kWalkChunk, lim_r);
    const bool have = cs < ce;
    // where the chain enters the round (kNoRel: the segment still guesses)
    const uint32_t pos_r = pos == kNone ? kNoRel : (uint32_t)min(pos - base, (uint64_t)0xFFFF0000u);

    // ---- pass A: guess and walk -------------------------------------------------------------
    WalkLane w;
    walk_clear(w);
    w.entry = kNoRel;
    w.exit = 0;
    if (have) {
      if (pos_r != kNoRel && pos_r >= cs && pos_r < ce) w.entry = pos_r;  // the exact entry lies in my chunk
      else if (pos_r != kNoRel && pos_r >= ce) w.entry = kNoRel;          // a record spans my whole chunk
      else w.entry = guess_entry(v, cs, ce, head_r);  // (the host's entry_hint is of no use here: it may
                                                      //  name a record that is not the FIRST one)
      if (w.entry != kNoRel) walk_chunk<AGG>(d, v, w.entry, ce, tmask, w);
    }

    // ---- seams: lane l must start where the records before it end --------------------------------
    uint32_t cur;  // where the chain leaves the round (kNoRel: still nothing found)
    {
      // if every guess is right: the chain position in front of my chunk is the exit of the nearest
      // walk before me (exits grow with the lane), else what the round inherited
      const uint32_t incl = warp_max_scan_incl(w.entry != kNoRel ? w.exit + 1u : 0u, lane);
      uint32_t excl = __shfl_up_sync(FULL, incl, 1);
      if (lane == 0) excl = 0;
      const uint32_t before = excl ? excl - 1 : pos_r;
      bool ok;
      if (before == kNoRel) ok = true;  // nothing known in front of me
      else if (!have || before >= ce) ok = w.entry == kNoRel;
      else ok = w.entry == before;
      ok = ok && w.err == 0;
      if (__all_sync(FULL, ok)) {
        const uint32_t last = __shfl_sync(FULL, incl, 31);
        cur = last ? last - 1 : pos_r;
      } else {
        // some guess was wrong (or a walk met an error): settle the lanes one after the other
        cur = pos_r;
        for (int l = 0; l < 32; ++l) {
          const uint32_t l_cs = max((uint32_t)l * kWalkChunk, lo_r);
          const uint32_t l_ce = min((uint32_t)(l + 1) * kWalkChunk, lim_r);
          if (l_cs >= l_ce) continue;
          if (cur == kNoRel) {
            // the segment's entry is still a guess: the first lane that found a start supplies it
            const uint32_t e = __shfl_sync(FULL, w.entry, l);
            if (e == kNoRel) continue;
            cur = e;
          }
          if (cur >= l_ce) {
            // no record starts in lane l's chunk: whatever it guessed was not a record
            if (lane == l) w.entry = kNoRel;
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
            err_pos = base + __shfl_sync(FULL, w.err_r, l);
            if ((int)lane > l) w.entry = kNoRel;  // records consume() never reaches
            break;
          }
        }
      }
    }
    if (entry == kNone) {
      const uint32_t m = __reduce_min_sync(FULL, w.entry);
      if (m != kNoRel) entry = base + m;
    }
    if (w.entry == kNoRel) walk_clear(w);
    const uint32_t stop = w.err ? w.err_r : ce;  // my records start in [w.entry, stop)

    // ---- pass B: outputs that need stream-order indices ------------------------------------------
    if (AGG == HBT_AGG_SWITCH_EVENTS) {
      const uint32_t incl = warp_sum_scan_incl(w.n_handled, lane);
      uint64_t row_i = n_rows + (incl - w.n_handled);
      if (w.n_handled) {
        uint32_t q = w.entry;
        while (q < stop) {
          const uint64_t h = v.near(q);
          if (counted(hdr_type(h), tmask)) {
            const ulonglong2 e = sw_proto_flat(v, q, h, d.tid_level, p.cpu_base + seg.cpu);
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
      const uint32_t incl_ev = warp_max_scan_incl(w.last_ev, lane);
      uint32_t in_rev = __shfl_up_sync(FULL, incl_ev, 1);
      if (lane == 0) in_rev = 0;
      const uint32_t in_kind = in_rev ? (in_rev & 3u) : ev_kind(ev);
      uint32_t my_rows, my_unknown;
      uint64_t my_seam;
      auto derive = [&]() {
        my_rows = w.n_rows;
        my_unknown = w.n_unknown;
        my_seam = 0;
        if ((w.first_ev & 3u) == EV_SAMPLE) {
          if (in_kind == EV_SAMPLE) ++my_rows;            // interval from the previous SAMPLE
          else if (in_kind == EV_LOST) ++my_unknown;      // re-seed, PerCpuCountSampleGenerator.h:135-154
          else my_seam = ev_make(base + rev_off(w.first_ev), EV_SAMPLE);  // nothing before it in this
        }                                                                  // segment: k_finalize's seam
      };
      derive();
      const uint32_t incl_r = warp_sum_scan_incl(my_rows, lane);
      const uint32_t row0 = incl_r - my_rows;  // my first row of the round
      int32_t late_err = 0;
      uint32_t late_r = 0;
      if (w.last_ev != 0) {
        // second walk: {current, previous} of each of my rows, checked like the generator checks them
        bool valid = in_kind == EV_SAMPLE;
        uint32_t prev = in_rev ? rev_off(in_rev) : kFarPrev;  // (kFarPrev: the SAMPLE ev names)
        uint32_t row_i = row0;
        uint32_t q = w.entry;
        while (q < stop) {
          const uint64_t h = v.near(q);
          const uint32_t type = hdr_type(h);
          if (counted(type, tmask)) {
            if (type == REC_SAMPLE && ev_samples) {
              if (valid) {
                const uint64_t ts = v.near(q + 8);
                const uint64_t pts = prev == kFarPrev ? ring_word_g(ring, ev_pos(ev) + 8) : v.near(prev + 8);
                if (pts > ts) {  // :156-163: the interval is refused, consume() stops in front of the record
                  late_err = -1;
                  late_r = q;
                  break;
                }
                if (n_rows + row_i >= seg.rows_cap) {  // (only with nr != n_events) a full output ring, :191-194
                  late_err = -28;
                  late_r = q;
                  break;
                }
                if (row_i < (uint32_t)kWalkMaxRows) sts_u32x2(s_desc + row_i * 8, q, prev);
                ++row_i;
              }
              valid = true;
              prev = q;
            } else if (type == REC_LOST) {
              valid = false;
            }
          }
          q += hdr_size(h);
        }
      }
      const uint32_t bad = __ballot_sync(FULL, late_err != 0);
      uint32_t round_rows = __shfl_sync(FULL, incl_r, 31);
      if (bad) {
        // consume() stops in front of that SAMPLE: what my walk counted behind it, and every lane
        // behind me, is dropped
        const int l = __ffs(bad) - 1;
        err = __shfl_sync(FULL, late_err, l);
        const uint32_t e_r = __shfl_sync(FULL, late_r, l);
        err_pos = base + e_r;
        if ((int)lane == l) {
          const uint32_t keep_entry = w.entry;
          walk_chunk<AGG>(d, v, keep_entry, e_r, tmask, w);
          derive();
          w.last_ev = rev_make(e_r, EV_LOST);  // count_.tstamp = invalid
        } else if ((int)lane > l) {
          w.entry = kNoRel;
          walk_clear(w);
          my_rows = my_unknown = 0;
          my_seam = 0;
        }
        round_rows = __reduce_add_sync(FULL, my_rows);
        cur = e_r;
      }
      __syncwarp();  // the descriptors are complete

      // rows of the round, written by the whole warp: lane -> one word of a row
      if (rows_per_iter) {
        for (uint32_t r0 = 0; r0 < round_rows; r0 += rows_per_iter) {
          const uint32_t rr = r0 + sub_row;
          if (sub_row < rows_per_iter && rr < round_rows) {
            const uint2 cp = lds_u32x2(s_desc + rr * 8);
            auto prev_word = [&](uint32_t off) {
              return cp.y == kFarPrev ? ring_word_g(ring, ev_pos(ev) + off) : v.far(cp.y + off);
            };
            uint64_t x;
            if (sub_col == 0) x = prev_word(8);
            else if (sub_col == 1) x = v.near(cp.x + 8) - prev_word(8);
            else {
              const uint32_t i = sub_col - 2;
              x = xform(d, i, v.far(cp.x + 40 + 8 * i)) - xform(d, i, prev_word(40 + 8 * i));
            }
            seg.rows[(1 + n_rows + rr) * row_words + sub_col] = x;
          }
        }
      } else {
        for (uint32_t rr = 0; rr < round_rows; ++rr) {
          const uint2 cp = lds_u32x2(s_desc + rr * 8);
          auto prev_word = [&](uint32_t off) {
            return cp.y == kFarPrev ? ring_word_g(ring, ev_pos(ev) + off) : v.far(cp.y + off);
          };
          for (uint32_t col = lane; col < row_words; col += 32) {
            uint64_t x;
            if (col == 0) x = prev_word(8);
            else if (col == 1) x = v.near(cp.x + 8) - prev_word(8);
            else {
              const uint32_t i = col - 2;
              x = xform(d, i, v.far(cp.x + 40 + 8 * i)) - xform(d, i, prev_word(40 + 8 * i));
            }
            seg.rows[(1 + n_rows + rr) * row_words + col] = x;
          }
        }
      }
      n_rows += round_rows;
      acc_unknown += my_unknown;
      const uint32_t le = __reduce_max_sync(FULL, w.last_ev);
      if (le) ev = ev_make(base + rev_off(le), le & 3u);
      const uint32_t ls = __reduce_max_sync(FULL, w.last_sample);
      if (ls) last_sample = base + ls - 1;
      if (seg_first_ev == 0) {
        const uint32_t seam_r = __reduce_max_sync(FULL, my_seam ? rev_off(w.first_ev) + 1 : 0u);
        if (seam_r) seg_first_ev = ev_make(base + seam_r - 1, EV_SAMPLE);
      }
    }

    // ---- keep the lanes' counts ------------------------------------------------------------------
    acc_rec += w.n_rec;
    acc_lost_n += w.n_lost;
    acc_lost += w.lost_sum;
    acc_lo += w.cnt_lo;
    acc_hi += w.cnt_hi;
    if (++acc_rounds == 7) flush();  // 7 rounds x 32 records per lane stay below 256
    pos = cur == kNoRel ? kNone : base + cur;  // exact from here on
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
