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
// 256-byte chunk of the round, guesses the first record start in it (first plausible header whose
// successor is plausible too) and then simply follows header.size from record to record -- work
// proportional to the RECORDS.  The seams between the 32 lanes are settled inside the warp: lane l's
// walk must begin where the records before it end; a lane whose guess was wrong walks again from the
// right place.  What leaves a round is exact, and the position after its last record carries into
// the next round, so only a segment's very first record remains a guess that k_finalize has to
// confirm (as for k_decode).  No block barriers: warps of a CTA share nothing.
//
// Round = stage; guess + walk (each lane marks the record starts of its chunk in a 32-bit mask, one
// bit per word); seam resolution; then the record starts of the whole round are written to one list
// in stream order and every per-record step -- type counters, size checks, proto-events, the
// SAMPLE / LOST interval state and the delta rows -- runs over that list 32 records at a time, one
// record per lane whatever its type or size, with ballots giving the stream-order indices.  The
// walk, whose trip count differs from lane to lane, is only a header load and an add per record.
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
constexpr int kWalkList = 512;                  // record starts listed at a time (16-bit offsets)
constexpr uint32_t kFarPrev = 0xFFFFFFFEu;      // "previous SAMPLE lies before this round"
static_assert(kWalkChunk == 256, "WalkView shifts by 8; a chunk's record starts fit a 32-bit mask");
static_assert(kWalkStageSmem % 8 == 0, "the lists follow the staging slice");

// per warp: staging slice, record-start list, COUNT_DELTAS: {current, previous} SAMPLE of up to 32 rows
__host__ __device__ constexpr int walk_smem_per_warp(int agg) {
  return kWalkStageSmem + kWalkList * 2 + (agg == HBT_AGG_COUNT_DELTAS ? 32 * 8 : 0);
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
    if (r < kWalkStagedBytes) return near(r);
    return slow(base + r);
  }
  __device__ __noinline__ uint64_t slow(uint64_t pos) const { return ring_word_g(*ring, pos); }
  // any ring position
  __device__ __forceinline__ uint64_t at(uint64_t pos) const {
    const uint64_t rel = pos - base;  // (positions before the round wrap to huge values)
    if (rel < (uint64_t)kWalkStagedBytes) return near((uint32_t)rel);
    return slow(pos);
  }
};

struct WalkLane {
  uint32_t entry;    // first record start in my chunk (kNoRel: none)
  uint32_t exit;     // where my walk left the chunk
  uint32_t mask;     // bit k: a record starts at word k of my chunk
  uint32_t bad;      // kNoRel, or the offset of a header that cannot be one (size < 8 or not a multiple of 8)
};

// Follow header.size from offset `from` while records start before `end` (both in one 256-byte
// aligned chunk), marking the starts.  Harmless on garbage (a guess may be wrong).
__device__ __forceinline__ void walk_chunk(const WalkView& v, uint32_t from, uint32_t end, WalkLane& w) {
  w.mask = 0;
  w.bad = kNoRel;
  uint32_t r = from;
  while (r < end) {
    const uint32_t size = (uint32_t)(v.near(r) >> 48);
    if (((size - 8u) & 0xFFFF0007u) != 0) {  // size < 8 or not a multiple of 8
      w.bad = r;
      break;
    }
    w.mask |= 1u << ((r >> 3) & 31u);  // word of my (256-byte aligned) chunk
    r += size;
  }
  w.exit = r;
}

// First record start at or after offset `from` (below `end`) judged by looks: a plausible header
// whose successor is plausible too (or lies beyond the staged bytes / the window).  kNoRel if there
// is none.  Only a guess: the seams decide.
__device__ __forceinline__ uint32_t guess_entry(const WalkView& v, uint32_t from, uint32_t end, uint32_t head_r) {
  for (uint32_t q = from; q < end; q += 8) {
    const uint64_t h = v.near(q);
    if (!plausible(h)) continue;
    const uint32_t nx = q + hdr_size(h);
    if (nx >= head_r || nx >= kWalkStagedBytes || plausible(v.near(nx))) return q;
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

// sw_min_size(type) / 8 for every type 0..15 (LOST 40, EXIT / FORK 48, everything else 32 bytes)
constexpr uint64_t kSwMinNib = 0x4444444444444444ull + (1ull << (4 * REC_LOST)) + (2ull << (4 * REC_EXIT)) +
                               (2ull << (4 * REC_FORK));

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

__device__ __forceinline__ void prefetch_l2(const void* p) {
  asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}
__device__ __forceinline__ void st_release_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void sts_u16(uint32_t saddr, uint32_t v) {
  asm volatile("st.shared.u16 [%0], %1;" ::"r"(saddr), "h"((unsigned short)v) : "memory");
}
__device__ __forceinline__ uint32_t lds_u16(uint32_t saddr) {
  unsigned short r;
  asm volatile("ld.shared.u16 %0, [%1];" : "=h"(r) : "r"(saddr));
  return r;
}

template <int AGG>
__global__ void __launch_bounds__(32 * kWalkWarps, 5) k_walk(const DecodeParams p, uint32_t n_launch) {
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
  const uint32_t lt = (1u << lane) - 1u;
  const bool group_layout = d.sample_layout == HBT_SAMPLE_GROUP;
  const bool ev_samples = AGG == HBT_AGG_COUNT_DELTAS && group_layout;
  const uint32_t row_words = 2 + d.n_events;
  // cooperative row writer: lane -> (row of the batch, word of the row)
  const uint32_t rows_per_iter = row_words <= 32 ? 32 / row_words : 0;
  const uint32_t sub_row = rows_per_iter ? lane / row_words : 0;
  const uint32_t sub_col = rows_per_iter ? lane - sub_row * row_words : lane;
  // word `col` of a row: column 0 = previous tstamp, 1 = tstamp - previous, 2 + i = value i - previous
  // value i (both through the TSC transform when bit i of the mask asks for it)
  auto col_off = [](uint32_t col) { return col < 2 ? 8u : 40u + 8u * (col - 2); };
  const uint32_t sub_off = col_off(sub_col);
  const bool sub_xf = sub_col >= 2 && ((d.tsc_transform_mask >> (sub_col - 2)) & 1);

  WalkView v;
  v.s = smem_u32(walk_smem) + warp * walk_smem_per_warp(AGG);
  v.ring = &ring;
  const uint32_t s_list = v.s + kWalkStageSmem;    // record starts of the round (16-bit offsets)
  const uint32_t s_desc = s_list + kWalkList * 2;  // COUNT_DELTAS: {current, previous} of up to 32 rows
  const bool src16 = (reinterpret_cast<uintptr_t>(ring.base) & 15) == 0;
  const uint32_t chunk = lane * kWalkChunk;

  // records are decoded while they START in [seg_lo, seg_limit)
  const uint64_t seg_lo = max(seg.tile_begin * (uint64_t)kTile, ring.tail);
  const uint64_t seg_limit = min(seg.tile_end * (uint64_t)kTile, ring.head);

  // Where the segment starts: exact for a CPU's first segment and for a forced re-run, else a guess
  // (kNone).  A guess is checked twice: here, against the exit of the segment in front as soon as that
  // one has published it (look-back below: a wrong guess is repaired by a second walk in the same
  // launch), and by k_finalize, which stays the authority and walks the whole chain afterwards.
  uint64_t start = seg.seg == 0 ? ring.tail : seg.forced_entry;
  bool published = false;
  for (uint32_t attempt = 0;; ++attempt) {
  // exact next record start, or kNone while the segment still guesses its entry
  uint64_t pos = start;
  uint64_t entry = kNone;  // first record start decoded (k_finalize checks it against the chain)
  int32_t err = 0;
  uint64_t err_pos = 0;
  // warp-uniform running state
  uint64_t n_rec = 0, n_rows = 0, n_unknown = 0;
  uint32_t n_lost = 0;
  uint64_t ev = 0;               // last SAMPLE / LOST event so far (ev_make)
  uint64_t last_sample = kNone;
  uint64_t seg_first_ev = 0;     // the segment's first SAMPLE when no event precedes it (the seam)
  // per-lane accumulators, reduced when the segment is done
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
    // the next round's lines on their way into L2 while this one is decoded
    if (base + kWalkRound < seg_limit) {
#pragma unroll
      for (uint32_t ln = lane; ln < kWalkStagedBytes / 128; ln += 32)
        prefetch_l2(ring.base + ((base + kWalkRound + ln * 128) & ring.mask));
    }

    // round-relative geometry: the segment's part of the round, my chunk's part of that
    const uint32_t lo_r = seg_lo > base ? (uint32_t)(seg_lo - base) : 0u;
    const uint32_t lim_r = (uint32_t)min(seg_limit - base, (uint64_t)kWalkRound);
    const uint32_t head_r = (uint32_t)min(ring.head - base, (uint64_t)0xFFFF0000u);
    const uint32_t cs = max(chunk, lo_r);
    const uint32_t ce = min(chunk + kWalkChunk, lim_r);
    const bool have = cs < ce;
    // where the chain enters the round (kNoRel: the segment still guesses)
    const uint32_t pos_r = pos == kNone ? kNoRel : (uint32_t)min(pos - base, (uint64_t)0xFFFF0000u);

    // ---- guess and walk -------------------------------------------------------------------------
    WalkLane w;
    w.entry = kNoRel;
    w.exit = 0;
    w.mask = 0;
    w.bad = kNoRel;
    if (have) {
      if (pos_r != kNoRel && pos_r >= cs && pos_r < ce) w.entry = pos_r;  // the exact entry lies in my chunk
      else if (pos_r != kNoRel && pos_r >= ce) w.entry = kNoRel;          // a record spans my whole chunk
      else w.entry = guess_entry(v, cs, ce, head_r);  // (the host's entry_hint is of no use here: it may
                                                      //  name a record that is not the FIRST one)
      if (w.entry != kNoRel) walk_chunk(v, w.entry, ce, w);
    }

    // ---- seams: lane l must start where the records before it end --------------------------------
    uint32_t cur;              // where the chain leaves the round (kNoRel: still nothing found)
    uint32_t bad_r = kNoRel;   // a header on the confirmed chain that cannot be one
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
      ok = ok && w.bad == kNoRel;
      if (__all_sync(FULL, ok)) {
        const uint32_t last = __shfl_sync(FULL, incl, 31);
        cur = last ? last - 1 : pos_r;
      } else {
        // some guess was wrong (or a walk met a bad header): settle the lanes one after the other
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
            walk_chunk(v, cur, ce, w);
          }
          cur = __shfl_sync(FULL, w.exit, l);
          bad_r = __shfl_sync(FULL, w.bad, l);
          if (bad_r != kNoRel) {  // consume() stops in front of that header (-EBADMSG) ...
            if ((int)lane > l) w.entry = kNoRel;  // ... and never reaches what lies behind it
            break;
          }
        }
      }
    }
    if (entry == kNone) {
      const uint32_t m = __reduce_min_sync(FULL, w.entry);
      if (m != kNoRel) entry = base + m;
    }
    if (w.entry == kNoRel) w.mask = 0;

    // ---- the round's records in stream order, 32 at a time -----------------------------------------
    const uint32_t my_n = __popc(w.mask);
    const uint32_t incl_n = warp_sum_scan_incl(my_n, lane);
    const uint32_t my_first = incl_n - my_n;
    const uint32_t n_round = __shfl_sync(FULL, incl_n, 31);
    for (uint32_t b0 = 0; b0 < n_round && err == 0; b0 += kWalkList) {
      __syncwarp();  // the readers of the previous list are done
      {
        uint32_t m = w.mask, k = my_first;
        while (m) {
          const uint32_t bit = __ffs(m) - 1;
          m &= m - 1;
          if (k - b0 < (uint32_t)kWalkList) sts_u16(s_list + (k - b0) * 2, chunk + bit * 8);
          ++k;
        }
      }
      __syncwarp();
      const uint32_t n_list = min(n_round - b0, (uint32_t)kWalkList);
      for (uint32_t i0 = 0; i0 < n_list && err == 0; i0 += 32) {
        bool act = i0 + lane < n_list;
        const uint32_t r = act ? lds_u16(s_list + (i0 + lane) * 2) : 0u;
        const uint64_t h = v.near(r);
        const uint32_t type = hdr_type(h), size = hdr_size(h);
        bool cnt = act && counted(type, tmask);
        // hard size checks of consume() (CpuEventsGroup.h:1435-1440, 1451-1456) and of the handlers
        int32_t e = 0;
        if (cnt && (type == REC_SAMPLE || type == REC_READ)) {
          const bool group = type == REC_READ || group_layout;
          const uint64_t want = group ? 40 + 8 * v.near(r + 16) : d.ip_sample_bytes;
          if ((uint64_t)size != want) e = -1;  // -EPERM
        }
        // (kSwMinNib: sw_min_size(type) / 8 in nibble `type`; the handler would read past a shorter record)
        if (AGG == HBT_AGG_SWITCH_EVENTS && cnt && size < 8u * ((uint32_t)(kSwMinNib >> (4 * (type & 15))) & 15u)) e = -1;
        // COUNT_DELTAS: the SAMPLE / LOST sequence (PerCpuCountSampleGenerator.h:122-195)
        bool is_s = false, is_l = cnt && type == REC_LOST, row = false, regress = false;
        uint32_t prev_kind = EV_NONE, prev_rel = kFarPrev, smask = 0, lmask = 0;
        if (AGG == HBT_AGG_COUNT_DELTAS) {
          is_s = cnt && e == 0 && type == REC_SAMPLE && ev_samples;
          smask = __ballot_sync(FULL, is_s);
          lmask = __ballot_sync(FULL, is_l);
          // the event in front of me: the nearest SAMPLE / LOST below my lane, else what came before
          const uint32_t below = (smask | lmask) & lt;
          const uint32_t pl = below ? 31 - __clz(below) : 0;
          const uint32_t r_pl = __shfl_sync(FULL, r, pl);
          if (below) {
            prev_kind = ((smask >> pl) & 1u) ? EV_SAMPLE : EV_LOST;
            prev_rel = r_pl;
          } else {
            prev_kind = ev_kind(ev);
            prev_rel = (ev != 0 && ev_pos(ev) >= base) ? (uint32_t)(ev_pos(ev) - base) : kFarPrev;
          }
          if (is_s && prev_kind == EV_SAMPLE) {
            const uint64_t ts = v.near(r + 8);
            const uint64_t pts = prev_rel == kFarPrev ? ring_word_g(ring, ev_pos(ev) + 8) : v.near(prev_rel + 8);
            if (pts > ts) {  // :156-163: the interval is refused, consume() stops in front of the record
              e = -1;
              regress = true;
            } else {
              row = true;
            }
          }
        }
        // consume() stops in front of the first record that fails: it and everything behind it is dropped
        bool invalidate = false;  // the failure came from the generator, which invalidates its interval
        auto cut = [&](uint32_t f, int32_t code) {
          err = code;
          err_pos = base + __shfl_sync(FULL, r, f);
          const bool keep = lane < f;
          act = act && keep;
          cnt = cnt && keep;
          is_s = is_s && keep;
          is_l = is_l && keep;
          row = row && keep;
          smask &= (1u << f) - 1u;
          lmask &= (1u << f) - 1u;
        };
        const uint32_t em = __ballot_sync(FULL, e != 0);
        if (em) {
          const uint32_t f = __ffs(em) - 1;
          invalidate = __shfl_sync(FULL, regress ? 1 : 0, f) != 0;
          cut(f, __shfl_sync(FULL, e, f));
        }
        uint32_t rowmask = 0;
        if (AGG == HBT_AGG_COUNT_DELTAS) {
          rowmask = __ballot_sync(FULL, row);
          // (only with nr != n_events records) the slab is a full output ring, :191-194
          const uint32_t full = __ballot_sync(FULL, row && n_rows + __popc(rowmask & lt) >= seg.rows_cap);
          if (full) {
            invalidate = true;
            cut(__ffs(full) - 1, -28);
            rowmask = __ballot_sync(FULL, row);
          }
        }

        // ---- counters ----------------------------------------------------------------------------
        n_rec += __popc(__ballot_sync(FULL, act));
        n_lost += __popc(__ballot_sync(FULL, is_l));
        if (is_l) acc_lost += v.near(r + 16);
        {
          const uint32_t ci = (uint32_t)(kCiNib >> (4 * (type & 15))) & 15u;
          const bool plain = cnt && !is_l;
          acc_lo += (plain && ci < 8) ? 1ull << (8 * (ci & 7)) : 0ull;
          acc_hi += (plain && ci >= 8) ? 1ull << (8 * (ci & 7)) : 0ull;
        }

        // ---- proto-events ------------------------------------------------------------------------
        if (AGG == HBT_AGG_SWITCH_EVENTS) {
          const uint32_t hm = __ballot_sync(FULL, cnt);
          if (cnt) {
            const uint64_t row_i = n_rows + __popc(hm & lt);
            const ulonglong2 pe = sw_proto_flat(v, r, h, d.tid_level, p.cpu_base + seg.cpu);
            if (row_i < seg.rows_cap) *reinterpret_cast<ulonglong2*>(seg.rows + (1 + row_i) * 2) = pe;
          }
          n_rows += __popc(hm);
        }

        // ---- delta rows --------------------------------------------------------------------------
        if (AGG == HBT_AGG_COUNT_DELTAS) {
          n_unknown += __popc(__ballot_sync(FULL, is_s && prev_kind == EV_LOST));  // re-seed, :135-154
          const uint32_t evm = smask | lmask;
          if (ev == 0 && evm != 0 && seg_first_ev == 0) {
            // the segment's first event: a SAMPLE here is the seam k_finalize joins to the segment before
            const uint32_t f = __ffs(evm) - 1;
            const uint32_t r_f = __shfl_sync(FULL, r, f);
            if ((smask >> f) & 1u) seg_first_ev = ev_make(base + r_f, EV_SAMPLE);
          }
          const uint32_t n_row = __popc(rowmask);
          if (n_row) {
            const uint64_t far_abs = ev_pos(ev);  // what kFarPrev names
            if (row) sts_u32x2(s_desc + __popc(rowmask & lt) * 8, r, prev_rel);
            __syncwarp();
            // written by the whole warp: lane -> one word of a row
            uint64_t* const out = seg.rows + (1 + n_rows) * row_words;
            auto emit = [&](uint32_t rr, uint32_t col, uint32_t off, bool xf) {
              const uint2 cp = lds_u32x2(s_desc + rr * 8);
              uint64_t b = cp.y == kFarPrev ? ring_word_g(ring, far_abs + off) : v.far(cp.y + off);
              uint64_t a = v.far(cp.x + off);
              if (xf) {
                a = tsc_to_ktime(a, d.tsc);
                b = tsc_to_ktime(b, d.tsc);
              }
              out[rr * row_words + col] = col == 0 ? b : a - b;
            };
            if (rows_per_iter) {
              for (uint32_t r0 = sub_row; r0 < n_row; r0 += rows_per_iter)
                if (sub_row < rows_per_iter) emit(r0, sub_col, sub_off, sub_xf);
            } else {
              for (uint32_t rr = 0; rr < n_row; ++rr)
                for (uint32_t col = lane; col < row_words; col += 32)
                  emit(rr, col, col_off(col), col >= 2 && ((d.tsc_transform_mask >> (col - 2)) & 1));
            }
            __syncwarp();  // the descriptors may be overwritten
            n_rows += n_row;
          }
          if (evm) {
            const uint32_t l = 31 - __clz(evm);
            ev = ev_make(base + __shfl_sync(FULL, r, l), ((smask >> l) & 1u) ? EV_SAMPLE : EV_LOST);
          }
          if (smask) last_sample = base + __shfl_sync(FULL, r, 31 - __clz(smask));
          if (invalidate) ev = ev_make(err_pos, EV_LOST);  // count_.tstamp = invalid
        }
      }
    }
    if (err == 0 && bad_r != kNoRel) {
      err = -74;  // -EBADMSG
      err_pos = base + bad_r;
    }
    if (err) cur = (uint32_t)(err_pos - base);
    if (++acc_rounds == 7) flush();  // 7 rounds x at most 32 records per lane stay below 256
    pos = cur == kNoRel ? kNone : base + cur;  // exact from here on
  }

  flush();
  const uint64_t lost_sum = warp_sum64(acc_lost);
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
    r.slow_tiles = attempt;  // 1: the segment was walked twice (its entry guess was wrong)
    r.hist_overflow = 0;
    r.n_lost_recs = n_lost;
  }
  if (p.seg_flag == nullptr) break;
  // ---- look-back ------------------------------------------------------------------------------
  // Publish my exit for the segment behind me -- at once when it can be trusted (the walk went through
  // to the segment's end: record chains that start apart meet again after a few records, so the exit
  // rarely depends on the entry), otherwise (the walk stopped at an error, maybe only because its
  // entry was wrong) after my own entry is settled.
  auto publish = [&]() {
    __syncwarp();
    if (lane == 0 && !published) {
      __threadfence();
      st_release_u32(p.seg_flag + seg_index, 1u);
    }
    published = true;
  };
  if (err == 0 || attempt != 0 || seg.seg == 0 || start != kNone) publish();
  if (attempt != 0 || seg.seg == 0 || start != kNone) break;  // my entry was exact, or is repaired
  // Compare my guess with the exit of the segment in front (same CPU: the table keeps a CPU's segments
  // together; it was launched before me, so it is running or done; the chain of waits ends at the
  // CPU's first segment, whose entry is exact).  The wait is bounded: should the flag never come,
  // k_finalize settles it the long way.
  uint64_t want = kNone;
  if (lane == 0) {
    uint32_t f = 0;
    for (uint32_t i = 0; i < (1u << 16) && !f; ++i) {
      f = ld_acquire_u32(p.seg_flag + seg_index - 1);
      if (!f) __nanosleep(400);
    }
    if (f) want = *reinterpret_cast<const volatile uint64_t*>(&p.results[seg_index - 1].exit);
  }
  want = __shfl_sync(FULL, want, 0);
  if (want == kNone || (want >= seg_limit ? entry == kNone : entry == want)) {  // no answer, or the guess was right
    publish();
    break;
  }
  start = want;                                                       // walk again from the right place
  }
}

}  // namespace hbt
