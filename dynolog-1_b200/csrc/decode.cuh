// Ring-buffer decode kernels (K1 boundary discovery, K2 dispatch + counters,
// K3 per-interval deltas, K6 (cpu,pid,ip) histogram) and the per-CPU finalize.
//
// Replaces CpuEventsGroup<TImpl,TMode>::consume
// (hbt/src/perf_event/CpuEventsGroup.h:1326-1522) and the handlers of
// CpuCountSampleGenerator (hbt/src/perf_event/PerCpuCountSampleGenerator.h:110-202)
// for all CPUs of a batch at once.
//
// Data layout in HBM.  One ring image per CPU (exactly the perf mmap data area,
// power-of-two bytes) and an unread window [tail, head) in absolute positions;
// physical offset = pos & (data_size-1), so the ring wrap (and the reference's
// two-piece memcpy, :1358-1370) disappears into the address computation.  The
// position space is cut into TILE-byte tiles aligned to TILE (a tile never
// straddles the ring end) and each CPU's tiles into `n_seg` contiguous segments.
// One CTA decodes one segment, tile after tile, with tiles staged in shared
// memory by 1-D bulk async copies (TMA unit) three tiles deep.
//
// Boundary discovery.  `tail += header.size` (:1503) is a serial chain.  Inside a
// tile every thread owns a 64..128-byte sub-chunk: it guesses the first record
// start in it (first plausible header whose successor is plausible too), walks
// its own records from there and reports where it left the sub-chunk.  A block
// wide max-scan then checks that every guess equals the exit of the records
// before it; by induction from the tile's exact entry the guesses are then the
// true boundaries.  A tile that fails the check (or holds any record the fast
// path does not want: bad size, timestamp regression, giant record) is decoded
// by one thread with a literal sequential consume() -- exactness never depends
// on the guess.  The same idea one level up: every segment but a CPU's first
// guesses its entry; k_finalize chains exits to entries and, on a mismatch,
// records the true entry so the host re-runs the batch with it forced.
#pragma once
#include <cstdint>
#include "../../include/hbt_b200.h"
#include "ptx.cuh"

namespace hbt {

constexpr uint32_t REC_LOST = 2, REC_COMM = 3, REC_EXIT = 4, REC_THROTTLE = 5,
                   REC_UNTHROTTLE = 6, REC_FORK = 7, REC_READ = 8, REC_SAMPLE = 9, REC_AUX = 11,
                   REC_ITRACE_START = 12, REC_SWITCH_CPU_WIDE = 15;

constexpr uint64_t kNone = ~0ull;          // "no position"
constexpr uint64_t kInvalidTs = ~0ull;     // kInvalidTimeStamp (System.h:97)
constexpr uint64_t kPrevCarry = ~0ull - 1; // "previous sample" = carry from the last call

#ifndef HBT_TILE_BYTES
#define HBT_TILE_BYTES 16384
#endif
#ifndef HBT_CTA_THREADS
#define HBT_CTA_THREADS 256
#endif
constexpr int kTile = HBT_TILE_BYTES;      // bytes per tile
constexpr int kStages = 3;                 // tiles resident per CTA
constexpr int kThreads = HBT_CTA_THREADS;  // threads per CTA
constexpr int kSubWords = kTile / 8 / kThreads;  // words (8 B) of the tile per thread
static_assert(kSubWords == 8 || kSubWords == 16, "skew formula below covers 8- and 16-word sub-chunks");

// Event kinds for the interval state machine of CpuCountSampleGenerator.
constexpr uint32_t EV_NONE = 0, EV_SAMPLE = 1, EV_LOST = 2;
__device__ __forceinline__ uint64_t ev_make(uint64_t pos, uint32_t kind) {
  return ((pos + 1) << 2) | kind;
}
__device__ __forceinline__ uint32_t ev_kind(uint64_t ev) { return (uint32_t)(ev & 3); }
__device__ __forceinline__ uint64_t ev_pos(uint64_t ev) { return (ev >> 2) - 1; }

struct RingDesc {  // one per CPU
  const uint8_t* base;
  uint64_t mask;      // data_size - 1
  uint64_t tail;
  uint64_t head;      // effective head (clamped on a re-run)
  uint64_t true_head; // head as submitted
  uint32_t seg_begin; // first segment id of this CPU
  uint32_t n_seg;
  int64_t override_ret;  // 0 = none; else forced consume() return on a re-run
  uint64_t* rows;        // row slab base for this CPU (COUNT_DELTAS)
};

struct SegDesc {  // one per CTA
  uint32_t cpu;
  uint32_t seg;          // index within the CPU
  uint64_t tile_begin;   // first tile index (absolute)
  uint64_t tile_end;     // one past the last tile index
  uint64_t forced_entry; // kNone = guess
  // Where the segment's first record starts IF the stream is one run of equal-sized records from
  // the window's tail (host arithmetic: tail + k * size); kNone = no such expectation.  Only a
  // first guess: the tile's headers must agree, and k_finalize still chains entry to exit.
  uint64_t entry_hint;
  uint32_t hint_size;    // that record size
  uint32_t pad_;
  uint64_t* rows;        // slab: slot 0 is the seam row, rows follow
  uint64_t rows_cap;     // rows after the seam slot
};

struct SegResult {
  uint64_t entry;  // first record start decoded by this segment, kNone if none
  uint64_t exit;   // start of the first record at or after the segment end, kNone if none
  uint64_t n_records;
  uint64_t counters[HBT_NUM_COUNTERS];
  uint64_t n_rows;
  uint64_t n_unknown;
  uint64_t first_ev;  // first SAMPLE/LOST event of the segment (encoded), 0 = none
  uint64_t last_ev;   // last SAMPLE/LOST event of the segment (encoded), 0 = none
  uint64_t last_sample;  // position of the last SAMPLE of the segment, kNone if none
  uint64_t err_pos;
  int32_t err;        // 0 or negative errno: decoding stopped before err_pos
  uint32_t slow_tiles;
  uint32_t hist_overflow;
  uint32_t n_lost_recs;  // handled LOST records of the segment (SWITCH_EVENTS: LOST proto-events)
};

struct CpuState {  // generator state across calls
  uint64_t counters[HBT_NUM_COUNTERS];
  uint64_t num_unknown;
  uint64_t carry_tstamp;
  uint64_t carry_last[HBT_MAX_EVENTS];
};

// (cpu,pid,ip) histogram storage: every CPU owns `buckets` append-only lists of
// 16-byte entries {u64 ip; u32 pid; u32 count}; a key always lands in the list
// its hash selects.  The decode kernel appends one entry (count 1) per sample;
// k_hist_aggregate then folds every list in shared memory into distinct keys,
// in place.  Lists accumulate over decodes until hbt_hist_clear.
struct HistLists {
  ulonglong2* entries;   // [n_cpus][buckets][cap]
  uint32_t* cursors;     // [n_cpus][buckets]: entries in each list
  uint32_t buckets;      // lists per CPU
  uint32_t cap;          // entries per list
  uint32_t* overflow;    // 1: a list ran out of room
};

struct DecodeParams {
  const RingDesc* rings;
  const SegDesc* segs;
  SegResult* results;
  const CpuState* state_in;
  hbt_decode_desc d;
  HistLists hist;
  uint32_t cpu_base;  // SWITCH_EVENTS: comp_unit_id = cpu_base + cpu
  // a re-run decodes only the segments listed here (null: CTA i decodes segment i)
  const uint32_t* run_list;
  // k_walk's look-back: per segment, 1 once its exit is in results[] (null: no look-back)
  uint32_t* seg_flag;
};

// ---------------------------------------------------------------------------
// Shared helpers
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t hdr_type(uint64_t h) { return (uint32_t)h; }
__device__ __forceinline__ uint32_t hdr_size(uint64_t h) { return (uint32_t)(h >> 48); }

// Guess filter only (never decides results): a header the kernel could have
// written -- small type id, u64-padded size.
__device__ __forceinline__ bool plausible(uint64_t h) {
  // type in [1, 64), size >= 8 and a multiple of 8 -- on the two halves of the header word
  // (size is the top 16 bits): three compares
  const uint32_t lo = (uint32_t)h, hi = (uint32_t)(h >> 32);
  return (lo - 1u) < 63u && (hi & 0x00070000u) == 0 && hi >= 0x00080000u;
}

__device__ __forceinline__ uint64_t ring_word_g(const RingDesc& r, uint64_t pos) {
  return *reinterpret_cast<const uint64_t*>(r.base + (pos & r.mask));
}

// CpuEventsGroup::kernelTimeFromTsc (CpuEventsGroup.h:1577-1591).
__device__ __forceinline__ uint64_t tsc_to_ktime(uint64_t tsc, const hbt_tsc_params& p) {
  const uint64_t rem_flags = (1ull << p.time_shift) - 1;
  const uint64_t quot = tsc >> p.time_shift;
  const uint64_t rem = tsc & rem_flags;
  return p.time_zero + quot * p.time_mult + ((rem * p.time_mult) >> p.time_shift);
}

__device__ __forceinline__ uint64_t xform(const hbt_decode_desc& d, uint32_t i, uint64_t v) {
  return ((d.tsc_transform_mask >> i) & 1) ? tsc_to_ktime(v, d.tsc) : v;
}

// Counter index of a record type, -1 for types consume() only skips
// (CpuEventsGroup.h:1372-1501).  Branch-free: a stream of mixed types would otherwise
// serialise a warp over a 16-way switch for every record.  type_mask(handled) has bit t
// set when records of PERF type t are counted; kCiNib holds the counter index of type t
// in nibble t.
__host__ __device__ constexpr uint32_t type_mask(uint32_t handled) {
  return ((handled & HBT_H_LOST) ? 1u << REC_LOST : 0u) | ((handled & HBT_H_COMM) ? 1u << REC_COMM : 0u) |
         ((handled & HBT_H_EXIT) ? 1u << REC_EXIT : 0u) | (1u << REC_THROTTLE) | (1u << REC_UNTHROTTLE) |
         ((handled & HBT_H_FORK) ? 1u << REC_FORK : 0u) | ((handled & HBT_H_SAMPLE) ? 1u << REC_SAMPLE : 0u) |
         ((handled & HBT_H_READ) ? 1u << REC_READ : 0u) | ((handled & HBT_H_AUX) ? 1u << REC_AUX : 0u) |
         ((handled & HBT_H_ITRACE_START) ? 1u << REC_ITRACE_START : 0u) |
         ((handled & HBT_H_SWITCH_CPU_WIDE) ? 1u << REC_SWITCH_CPU_WIDE : 0u);
}
constexpr uint64_t kCiNib =
    ((uint64_t)HBT_CNT_LOST << (4 * REC_LOST)) | ((uint64_t)HBT_CNT_COMM << (4 * REC_COMM)) |
    ((uint64_t)HBT_CNT_EXIT << (4 * REC_EXIT)) | ((uint64_t)HBT_CNT_THROTTLE << (4 * REC_THROTTLE)) |
    ((uint64_t)HBT_CNT_UNTHROTTLE << (4 * REC_UNTHROTTLE)) | ((uint64_t)HBT_CNT_FORK << (4 * REC_FORK)) |
    ((uint64_t)HBT_CNT_SAMPLE << (4 * REC_SAMPLE)) | ((uint64_t)HBT_CNT_READ << (4 * REC_READ)) |
    ((uint64_t)HBT_CNT_AUX << (4 * REC_AUX)) | ((uint64_t)HBT_CNT_ITRACE_START << (4 * REC_ITRACE_START)) |
    ((uint64_t)HBT_CNT_SWITCH_CPU_WIDE << (4 * REC_SWITCH_CPU_WIDE));
static_assert(REC_SWITCH_CPU_WIDE == 15 && HBT_NUM_COUNTERS <= 16, "nibble table covers types 0..15");

__device__ __forceinline__ bool counted(uint32_t type, uint32_t tmask) {
  return type < 16 && ((tmask >> (type & 15)) & 1u);
}
__device__ __forceinline__ int counter_of(uint32_t type, uint32_t tmask) {
  return counted(type, tmask) ? (int)((kCiNib >> (4 * (type & 15))) & 15) : -1;
}

// ---- HBT_AGG_SWITCH_EVENTS ----------------------------------------------------------
// Pass 1 (k_decode) writes one 16-byte "proto-event" per handled mode::ContextSwitch
// record into the segment's row slab, in stream order: the tagstack::Event the record's
// handler emits (PerCpuThreadSwitchGenerator.h:228-412) with the record's own timestamp.
// What depends on the PREVIOUS record -- the WriteErrorsEnd in front of the first record
// after a LOST (pendingRecordLost_, :441-453) and the WriteErrorsStart timestamp
// last_tstamp_ + 1 (:243-245) -- is resolved by pass 2 (k_sw_scan / k_sw_emit),
// which is a prefix sum over "1 + previous is LOST".
constexpr uint32_t kMiscSwitchOut = 1u << 13, kMiscSwitchOutPreempt = 1u << 14;

// smallest well-formed size of a handled ContextSwitch record (CpuEventsGroup.h:105-184);
// shorter ones are refused with -EPERM (the reference would read past the record)
__device__ __forceinline__ uint32_t sw_min_size(uint32_t type) {
  // LOST 40, EXIT / FORK 48, COMM / SWITCH_CPU_WIDE 32 (selects, no branches)
  return 32u + (type == REC_LOST ? 8u : 0u) + ((type == REC_EXIT || type == REC_FORK) ? 16u : 0u);
}

__device__ __forceinline__ uint64_t sw_word1(uint32_t type, uint32_t level, uint32_t cu, uint32_t tag) {
  return (uint64_t)(type & 0xFF) | ((uint64_t)(level & 0xFF) << 8) | ((uint64_t)(cu & 0xFFFF) << 16) |
         ((uint64_t)tag << 32);
}

// Proto-event of the handled record whose header is h; at(off) reads the record's word at
// byte offset off.  cu = comp_unit_id.
template <class At>
__device__ __forceinline__ ulonglong2 sw_proto(uint64_t h, uint32_t level, uint32_t cu, At at) {
  const uint32_t type = hdr_type(h), size = hdr_size(h);
  switch (type) {
    case REC_LOST:  // hdr id num_lost {pid tid tstamp}
      return make_ulonglong2(at(32), sw_word1(HBT_SWEV_WRITE_ERRORS_START, HBT_LEVEL_INVALID, cu, HBT_TAG_NA));
    case REC_COMM:  // hdr pid tid comm[] ... {pid tid tstamp}
      return make_ulonglong2(at(size - 8), sw_word1(HBT_SWEV_COMM, level, cu, (uint32_t)(at(8) >> 32)));
    case REC_EXIT:  // hdr pid ppid tid ptid time {..}
      return make_ulonglong2(at(24), sw_word1(HBT_SWEV_THREAD_DESTRUCTION, level, cu, (uint32_t)at(16)));
    case REC_FORK:
      return make_ulonglong2(at(24), sw_word1(HBT_SWEV_THREAD_CREATION, level, cu, (uint32_t)at(16)));
    default: {      // SWITCH_CPU_WIDE: hdr next_prev_pid next_prev_tid {pid tid tstamp}
      const uint32_t misc = (uint32_t)(h >> 32) & 0xFFFF;
      if (misc & kMiscSwitchOut)
        return make_ulonglong2(at(24), sw_word1((misc & kMiscSwitchOutPreempt) ? HBT_SWEV_SWITCH_OUT_PREEMPT
                                                                             : HBT_SWEV_SWITCH_OUT_YIELD,
                                                level, cu, HBT_TAG_NA));
      return make_ulonglong2(at(24), sw_word1(HBT_SWEV_SWITCH_IN, level, cu, (uint32_t)(at(16) >> 32)));
    }
  }
}

__device__ __forceinline__ uint32_t hist_hash(uint64_t ip, uint32_t pid) {
  // 32-bit multiply-mix (murmur3 finaliser): the top bits choose the list, the low bits
  // the cache set, all of it the cache tag.  64-bit multiplies would cost three times
  // as many instructions on the per-sample path.
  uint32_t x = (uint32_t)ip * 0xCC9E2D51u + (uint32_t)(ip >> 32) * 0x1B873593u + pid * 0xE6546B64u;
  x ^= x >> 16;
  x *= 0x85EBCA6Bu;
  x ^= x >> 13;
  x *= 0xC2B2AE35u;
  x ^= x >> 16;
  return x;
}

// Append one sample to the list of its bucket: one atomic on a cursor that lives
// in L2 (the cursors of the CPUs in flight are a few KB) and one 16-byte store.
// Neighbouring entries of a list are written close in time, so L2 merges them
// into full sectors before they reach HBM.
__device__ __forceinline__ uint32_t hist_reserve(const HistLists& t, uint32_t cpu, uint32_t hash,
                                                 uint32_t& list) {
  const uint32_t b = (uint32_t)(((uint64_t)hash * t.buckets) >> 32);
  list = cpu * t.buckets + b;
  return atomicAdd(t.cursors + list, 1u);
}
__device__ __forceinline__ void hist_store(const HistLists& t, uint32_t list, uint32_t pos,
                                           uint64_t ip, uint32_t pid, uint32_t weight) {
  if (pos < t.cap) {
    t.entries[(uint64_t)list * t.cap + pos] =
        make_ulonglong2(ip, (uint64_t)pid | ((uint64_t)weight << 32));
  } else {
    atomicExch(t.overflow, 1u);
  }
}
__device__ __forceinline__ void hist_append(const HistLists& t, uint32_t cpu, uint64_t ip,
                                            uint32_t pid, uint32_t weight, uint32_t hash) {
  uint32_t list;
  const uint32_t pos = hist_reserve(t, cpu, hash, list);
  hist_store(t, list, pos, ip, pid, weight);
}

// Shared-memory privatised (pid,ip) cache of one segment.  Its job is the heavy
// hitters: a key that occurs thousands of times in a CPU's window would otherwise
// append thousands of entries to one list.  Two-way set associative, 16-byte slots
// {ip, pid | count << 32}; an all-zero slot is empty (a live slot has count >= 1).
// A slot is claimed AND counted by one 128-bit shared-memory CAS, so there is never
// a half-written slot to wait for.  A 32-bit tag per slot (0 = never claimed) is only
// a filter: most samples are not in the cache, and one 8-byte load of the set's two
// tags tells them so; a tag match is always confirmed against the full key.  A sample
// whose set holds other keys is appended to its list directly.  Every sample is
// counted exactly once.  Flushed to the lists (weighted entries) at segment end.
#ifndef HBT_CACHE_SLOTS
#define HBT_CACHE_SLOTS 1024
#endif
constexpr int kCacheSlots = HBT_CACHE_SLOTS;

struct HistCache {
  ulonglong2 slots[kCacheSlots];
  uint32_t tags[kCacheSlots];
};

// Count the sample in the segment cache; false = no room for this key (append it).
__device__ __forceinline__ bool hist_cache_count(HistCache& c, uint64_t ip, uint32_t pid,
                                                 uint32_t& hash) {
  hash = hist_hash(ip, pid);
  const uint32_t set = (hash ^ (hash >> 16)) & (kCacheSlots - 2);  // slots set, set + 1
  const uint32_t tag = hash | 1u;
  const uint64_t tg = *reinterpret_cast<const volatile uint64_t*>(c.tags + set);
#pragma unroll
  for (int way = 0; way < 2; ++way) {
    const uint32_t t = way == 0 ? (uint32_t)tg : (uint32_t)(tg >> 32);
    const uint32_t i = set + way;
    if (t == tag) {
      const volatile ulonglong2* vs = c.slots + i;
      const uint64_t lo = vs->x, hi = vs->y;
      if (lo == ip && (uint32_t)hi == pid && (hi >> 32) != 0) {
        atomicAdd(reinterpret_cast<uint32_t*>(c.slots + i) + 3, 1u);
        return true;
      }
    } else if (t == 0) {
      uint64_t lo, hi;
      cas128_shared(smem_u32(c.slots + i), 0, 0, ip, (uint64_t)pid | (1ull << 32), lo, hi);
      if (lo == 0 && hi == 0) {  // claimed with count 1
        *reinterpret_cast<volatile uint32_t*>(c.tags + i) = tag;
        return true;
      }
      if (lo == ip && (uint32_t)hi == pid) {  // someone claimed it for the same key just now
        atomicAdd(reinterpret_cast<uint32_t*>(c.slots + i) + 3, 1u);
        return true;
      }
    }
  }
  return false;
}

// Up to two appends of one thread whose slot reservations (hist_reserve) have been issued
// but whose entries are not stored yet.
struct HistPending {
  uint64_t ip0, ip1;
  uint32_t pid0, pid1, l0, l1, q0, q1;
  bool m0 = false, m1 = false;
  __device__ __forceinline__ void flush(const HistLists& t) {
    if (m0) hist_store(t, l0, q0, ip0, pid0, 1);
    if (m1) hist_store(t, l1, q1, ip1, pid1, 1);
    m0 = m1 = false;
  }
};

__device__ __forceinline__ void hist_count(const HistLists& t, HistCache& c, uint32_t cpu,
                                           uint64_t ip, uint32_t pid) {
  uint32_t hash;
  if (!hist_cache_count(c, ip, pid, hash)) hist_append(t, cpu, ip, pid, 1, hash);
}

// ---------------------------------------------------------------------------
// Per-CTA decode state (shared memory)
// ---------------------------------------------------------------------------
struct SegState {
  uint64_t pos;       // next record start; kNone while a segment still guesses
  uint64_t ev;        // last SAMPLE/LOST event so far in this segment (encoded)
  uint64_t first_ev;  // first SAMPLE of the segment when nothing precedes it (the seam)
  uint64_t last_sample;  // position of the last SAMPLE so far, kNone if none
  uint64_t entry;     // first record start decoded
  unsigned long long n_rec, n_rows, n_unknown;
  unsigned long long counters[HBT_NUM_COUNTERS];
  uint64_t err_pos;
  int32_t err;
  uint32_t slow_tiles;
  uint32_t n_lost_recs;  // handled LOST records so far
  // per-tile scratch, one copy per guessing attempt (0: stride-predicted, 1: scanned)
  uint32_t guess_min[2];  // smallest guessed entry (tile-relative), segment start only
  uint32_t tile_x[2];     // where the record chain left the tile (tile-relative)
  uint32_t size_min, size_max;  // record sizes seen in the tile
  uint32_t stride;    // record size when the previous tile held one size only, else 0
};

constexpr uint32_t kNoRel = 0xFFFFFFFFu;  // "no tile-relative position"

struct Smem {
  uint64_t bar[kStages];
  SegState st;
  uint32_t expect[kThreads];  // expect[j]: a record was seen to start at this offset of sub-chunk j
  uint32_t warp_ev[kThreads / 32];
  uint32_t warp_rows[kThreads / 32];
  uint32_t warp_ls[kThreads / 32];  // last SAMPLE start in the tile, relative + 1
};

// COUNT_DELTAS: (current SAMPLE, previous SAMPLE) positions of every row of the tile
// in row order, so that the rows can be written by consecutive threads.
constexpr int kMaxRowsPerTile = kTile / 48 + 1;  // a SAMPLE is at least 40 + 8 bytes
struct RowDescs {
  uint64_t far_prev;            // absolute position of a previous SAMPLE outside the tile
  uint2 d[kMaxRowsPerTile];     // {current, previous} tile-relative; previous = kNoRel -> far_prev
  // The last SAMPLE of a uniform tile, kept for the first row of the next tile (its bytes have
  // left the staging area by then; without this copy that row waits for global memory).
  // Two slots by tile parity: tile k writes slot k & 1 while its own first row reads the other.
  uint64_t prev_pos[2];                          // absolute position of the saved SAMPLE, kNone = none
  uint64_t prev_words[2][5 + HBT_MAX_EVENTS];    // its header, tstamp, nr, times and values
};

// Sequential consume() over [st.pos, limit) by ONE thread: the slow path for
// tiles the fast path declines and the definition of what the fast path must
// equal.  Reads the ring through global memory.  Stops at `limit` (first record
// start >= limit is left in st.pos) or at an error.
template <int AGG>
__device__ void consume_sequential(const DecodeParams& p, const RingDesc& ring,
                                   const SegDesc& seg, SegState& st, uint64_t limit) {
  const hbt_decode_desc& d = p.d;
  uint64_t pos = st.pos;
  // record sizes seen by this walk: a tile of one record size re-arms the stride prediction, so one
  // tile the guesses could not handle does not send the rest of its segment down this path as well
  st.size_min = kNoRel;
  st.size_max = 0;
  while (pos < limit) {
    const uint64_t h = ring_word_g(ring, pos);
    const uint32_t type = hdr_type(h), size = hdr_size(h);
    if (size < 8 || (size & 7)) {
      st.err = -74;  // -EBADMSG
      st.err_pos = pos;
      break;
    }
    st.size_min = min(st.size_min, size);
    st.size_max = max(st.size_max, size);
    if (st.entry == kNone) st.entry = pos;
    const int ci = counter_of(type, type_mask(d.handled_mask));
    if (type == REC_LOST && ci >= 0) {
      ++st.n_lost_recs;
      st.counters[HBT_CNT_LOST] += ring_word_g(ring, pos + 16);
      if (AGG == HBT_AGG_COUNT_DELTAS) st.ev = ev_make(pos, EV_LOST);
    } else if (type == REC_SAMPLE && ci >= 0) {
      if (d.sample_layout == HBT_SAMPLE_GROUP) {
        const uint64_t nr = ring_word_g(ring, pos + 16);
        if ((uint64_t)size != 40 + 8 * nr) {
          st.err = -1;  // -EPERM
          st.err_pos = pos;
          break;
        }
        if (AGG == HBT_AGG_COUNT_DELTAS) {
          const uint64_t ts = ring_word_g(ring, pos + 8);
          const uint32_t k = ev_kind(st.ev);
          if (k == EV_NONE) {
            // nothing before in this segment: the seam with the previous segment
            // (or the carry of the previous call) is resolved by k_finalize.
            st.first_ev = ev_make(pos, EV_SAMPLE);
          } else if (k == EV_LOST) {
            ++st.n_unknown;  // re-seed, PerCpuCountSampleGenerator.h:135-154
          } else {
            const uint64_t ppos = ev_pos(st.ev);
            const uint64_t pts = ring_word_g(ring, ppos + 8);
            if (pts > ts) {  // :156-163
              st.ev = ev_make(pos, EV_LOST);  // count_.tstamp = invalid
              st.err = -1;
              st.err_pos = pos;
              break;
            }
            if (st.n_rows >= seg.rows_cap) {
              // only reachable with nr != n_events records (the reference DCHECKs
              // it, :133): behaves like a full output ring, :191-194
              st.ev = ev_make(pos, EV_LOST);
              st.err = -28;  // -ENOSPC
              st.err_pos = pos;
              break;
            }
            uint64_t* row = seg.rows + (1 + st.n_rows) * (2 + (uint64_t)d.n_events);
            row[0] = pts;
            row[1] = ts - pts;
            for (uint32_t i = 0; i < d.n_events; ++i) {
              const uint64_t v = xform(d, i, ring_word_g(ring, pos + 40 + 8ull * i));
              const uint64_t pv = xform(d, i, ring_word_g(ring, ppos + 40 + 8ull * i));
              row[2 + i] = v - pv;
            }
            ++st.n_rows;
          }
          st.ev = ev_make(pos, EV_SAMPLE);
          st.last_sample = pos;
        }
      } else {
        if (size != d.ip_sample_bytes) {
          st.err = -1;
          st.err_pos = pos;
          break;
        }
        if (AGG == HBT_AGG_IP_HISTOGRAM) {
          const uint64_t ip = ring_word_g(ring, pos + 8);
          const uint32_t pid = (uint32_t)ring_word_g(ring, pos + 16);
          hist_append(p.hist, seg.cpu, ip, pid, 1, hist_hash(ip, pid));
        }
      }
      ++st.counters[HBT_CNT_SAMPLE];
    } else if (type == REC_READ && ci >= 0) {
      const uint64_t nr = ring_word_g(ring, pos + 16);
      if ((uint64_t)size != 40 + 8 * nr) {
        st.err = -1;
        st.err_pos = pos;
        break;
      }
      ++st.counters[HBT_CNT_READ];
    } else if (ci >= 0) {
      ++st.counters[ci];
    }
    if (AGG == HBT_AGG_SWITCH_EVENTS && ci >= 0) {
      if (size < sw_min_size(type) || st.n_rows >= seg.rows_cap) {
        // take the counter update back: the handler fails before consume() counts the record
        if (type == REC_LOST) {
          st.counters[HBT_CNT_LOST] -= ring_word_g(ring, pos + 16);
          --st.n_lost_recs;
        } else {
          --st.counters[ci];
        }
        st.err = size < sw_min_size(type) ? -1 : -28;  // -EPERM / -ENOSPC
        st.err_pos = pos;
        break;
      }
      const ulonglong2 e = sw_proto(h, d.tid_level, p.cpu_base + seg.cpu,
                                    [&](uint32_t off) { return ring_word_g(ring, pos + off); });
      uint64_t* row = seg.rows + (1 + st.n_rows) * 2;
      row[0] = e.x;
      row[1] = e.y;
      ++st.n_rows;
    }
    ++st.n_rec;
    pos += size;
  }
  st.pos = pos;
}

// The staged bytes of tile `t` (stage a) and tile t+1 (stage b) addressed by
// 32-bit tile-relative offsets: all hot-loop arithmetic is 32-bit.
struct TileView {
  uint32_t sa;  // shared-space address of tile t
  uint32_t sb;  // shared-space address of tile t+1, minus kTile
  uint64_t t_begin;
  const RingDesc* ring;
  // rel in [0, 2*kTile)
  __device__ __forceinline__ uint64_t at(uint32_t rel) const {
    return lds_u64((rel < (uint32_t)kTile ? sa : sb) + rel);
  }
  // any absolute position (older tiles / carry live in global memory)
  __device__ __forceinline__ uint64_t word_abs(uint64_t pos) const {
    const uint64_t off = pos - t_begin;  // wraps for pos < t_begin -> huge
    if (off < 2ull * kTile) return at((uint32_t)off);
    return ring_word_g(*ring, pos);
  }
};

// First word (tile-relative, in 8-byte words) of sub-chunk j.  The skew spreads
// the lock-step header scan of a warp over all shared-memory banks.
__device__ __forceinline__ uint32_t sub_begin_word(uint32_t j) {
  // a warp's 32 sub-chunk starts must cover the 16 eight-byte bank pairs twice
  return kSubWords * j + (kSubWords == 8 ? ((j >> 1) & 7) : (j & 15));
}
// Sub-chunk that holds tile word w.
__device__ __forceinline__ uint32_t sub_of_word(uint32_t w) {
  const uint32_t j = w / kSubWords;
  return w < sub_begin_word(j) ? j - 1 : j;
}

template <int AGG>
#ifndef HBT_MIN_BLOCKS
#define HBT_MIN_BLOCKS 4       // CTAs per SM the register budget must allow
#endif
#ifndef HBT_HIST_MIN_BLOCKS
#define HBT_HIST_MIN_BLOCKS 3  // IP_HISTOGRAM: the segment cache leaves room for three
#endif
__global__ void __launch_bounds__(kThreads, AGG == HBT_AGG_IP_HISTOGRAM ? HBT_HIST_MIN_BLOCKS : HBT_MIN_BLOCKS) k_decode(const DecodeParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* stages = smem_raw;  // kStages * kTile
  Smem& sm = *reinterpret_cast<Smem*>(smem_raw + kStages * kTile);
  HistCache& cache =
      *reinterpret_cast<HistCache*>(smem_raw + kStages * kTile + ((sizeof(Smem) + 15) & ~15ull));
  RowDescs& rowd =
      *reinterpret_cast<RowDescs*>(smem_raw + kStages * kTile + ((sizeof(Smem) + 15) & ~15ull));

  const uint32_t seg_index = p.run_list ? p.run_list[blockIdx.x] : blockIdx.x;
  const SegDesc seg = p.segs[seg_index];
  const RingDesc ring = p.rings[seg.cpu];
  const hbt_decode_desc& d = p.d;
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  SegState& st = sm.st;
  const bool h_sample = (d.handled_mask & HBT_H_SAMPLE) != 0;
  const bool h_read = (d.handled_mask & HBT_H_READ) != 0;
  const bool h_lost = (d.handled_mask & HBT_H_LOST) != 0;
  const bool group_layout = d.sample_layout == HBT_SAMPLE_GROUP;
  const uint32_t tmask = type_mask(d.handled_mask);

  if (tid == 0) {
    for (int s = 0; s < kStages; ++s) mbar_init(&sm.bar[s], 1);
    fence_mbar_init();
    st.pos = seg.seg == 0 ? ring.tail : seg.forced_entry;
    st.ev = 0;
    st.first_ev = 0;
    st.last_sample = kNone;
    st.entry = kNone;
    st.n_rec = st.n_rows = st.n_unknown = 0;
    for (int i = 0; i < HBT_NUM_COUNTERS; ++i) st.counters[i] = 0;
    st.err = 0;
    st.err_pos = 0;
    st.slow_tiles = 0;
    st.n_lost_recs = 0;
    st.guess_min[0] = st.guess_min[1] = kNoRel;
    st.tile_x[0] = st.tile_x[1] = kNoRel;
    st.stride = 0;
    st.size_min = kNoRel;
    st.size_max = 0;
  }
  sm.expect[tid] = kNoRel;
  if (AGG == HBT_AGG_COUNT_DELTAS && tid < 2) rowd.prev_pos[tid] = kNone;
  if (AGG == HBT_AGG_IP_HISTOGRAM) {
    for (int i = tid; i < kCacheSlots; i += kThreads) {
      cache.slots[i] = make_ulonglong2(0, 0);
      cache.tags[i] = 0;
    }
  }
  __syncthreads();

  const uint32_t n_tiles = (uint32_t)(seg.tile_end - seg.tile_begin);
  // Records are decoded while they START before seg_limit.
  const uint64_t seg_limit = min(seg.tile_end * (uint64_t)kTile, ring.head);
  const uint64_t policy = l2_policy_evict_first();

  // shared-space addresses, computed once and kept (see pin_u32)
  uint32_t stages_u32 = smem_u32(stages), bar_u32 = smem_u32(&sm.bar[0]);
  pin_u32(stages_u32);
  pin_u32(bar_u32);
  auto issue = [&](uint32_t k, uint32_t stage) {  // tile k of this segment (k == n_tiles: lookahead)
    const uint64_t t = seg.tile_begin + k;
    mbar_arrive_expect_tx_u32(bar_u32 + 8 * stage, kTile);
    bulk_g2s_hint_u32(stages_u32 + stage * kTile, ring.base + ((t * (uint64_t)kTile) & ring.mask), kTile,
                      bar_u32 + 8 * stage, policy);
  };
  // stage of tile k is k % kStages (tile numbering local to the segment)
  if (tid == 0) {
    issue(0, 0);
    issue(1, 1);
  }
  uint32_t phase_bits = 0;  // mbarrier parity per stage
  uint32_t sa_i = 0;        // stage of the current tile

  // thread-local totals over the fast-path tiles (added to st in the epilogue)
  uint32_t my_rec = 0, my_sample = 0;
  // per-type counts of the other record types (all but LOST and SAMPLE: nine counters), 7 bits
  // each in ONE register: counter ci lives at bit 7 * other_field(ci); flushed to shared memory
  // every 127 records.  One shared-memory atomic per record would serialise the whole CTA on a
  // stream of SWITCH records, and an indexed array of registers would live in local memory.
  uint64_t my_other = 0;
  uint32_t my_other_n = 0;
  // LOST: sum of num_lost and number of records, added to st in the epilogue
  uint64_t my_lost_sum = 0;
  uint32_t my_lost_n = 0;
  auto other_field = [](int ci) { return ci - 1 - (ci > HBT_CNT_SAMPLE ? 1 : 0); };
  auto flush_other = [&](bool warp_wide) {
#pragma unroll
    for (int c = 0; c < HBT_NUM_COUNTERS; ++c) {
      if (c == HBT_CNT_LOST || c == HBT_CNT_SAMPLE) continue;
      uint32_t v = (uint32_t)(my_other >> (7 * other_field(c))) & 0x7Fu;
      if (warp_wide) {
        v = __reduce_add_sync(0xffffffffu, v);
        if (lane != 0) v = 0;
      }
      if (v) atomicAdd(&st.counters[c], (unsigned long long)v);
    }
    my_other = 0;
  };

  HistPending pend;  // IP_HISTOGRAM: appends whose cursor atomics are in flight
  // COUNT_DELTAS row writer (uniform fast path): this thread's fixed word column and first row
  const uint32_t rw_wpr = 2 + d.n_events;
  // {first row, word column, rows per pass, active}: packed and pinned, or the compiler
  // re-derives them with two integer divisions in every tile
  uint32_t rw_pack;
  {
    const uint32_t rows = kThreads / rw_wpr;  // whole rows per pass (n_events <= 64: >= 3)
    const uint32_t r0 = tid / rw_wpr;
    rw_pack = r0 | ((tid - r0 * rw_wpr) << 8) | (rows << 16) | ((tid < rows * rw_wpr ? 1u : 0u) << 24);
    pin_u32(rw_pack);
  }
  const uint32_t rw_r0 = rw_pack & 0xFF, rw_w = (rw_pack >> 8) & 0xFF, rw_rows = (rw_pack >> 16) & 0xFF;
  const bool rw_active = (rw_pack >> 24) != 0;

  // my sub-chunk, in tile-relative bytes
  const uint32_t sb = 8u * sub_begin_word(tid);
  const uint32_t se_full = tid == kThreads - 1 ? (uint32_t)kTile : 8u * sub_begin_word(tid + 1);

#pragma unroll 1
  for (uint32_t k = 0; k < n_tiles; ++k) {
    const uint32_t tile_k = k;  // (inner loops reuse the name k)
    const uint32_t sb_i = sa_i + 1 == kStages ? 0 : sa_i + 1;
    const uint32_t sc_i = sb_i + 1 == kStages ? 0 : sb_i + 1;
    // refill the stage tile k-1 lived in; the trailing barrier of the previous
    // round ordered every read of it before this copy
    if (tid == 0 && k + 2 <= n_tiles) issue(k + 2, sc_i);
    if (k == 0) {
      mbar_wait_u32(bar_u32 + 8 * sa_i, (phase_bits >> sa_i) & 1);
      phase_bits ^= 1u << sa_i;
    }
    mbar_wait_u32(bar_u32 + 8 * sb_i, (phase_bits >> sb_i) & 1);
    phase_bits ^= 1u << sb_i;

    const uint64_t t_begin = (seg.tile_begin + k) * (uint64_t)kTile;
    TileView tv{stages_u32 + sa_i * (uint32_t)kTile, stages_u32 + sb_i * (uint32_t)kTile - (uint32_t)kTile,
                t_begin, &ring};
    // tile-uniform limits, tile-relative: record starts must be < lim; bytes up to
    // look (<= 2*kTile) are resident and inside the window
    const uint32_t lim =
        seg_limit > t_begin ? (uint32_t)min((uint64_t)kTile, seg_limit - t_begin) : 0u;
    const uint32_t look =
        ring.head > t_begin ? (uint32_t)min((uint64_t)(2 * kTile), ring.head - t_begin) : 0u;
    const uint64_t entry_abs = st.pos;  // exact, or kNone while the segment still guesses
    const bool stopped = st.err != 0;
    // entry relative to the tile: kNoRel = unknown, >= kTile = beyond this tile
    uint32_t erel;
    if (entry_abs == kNone) erel = kNoRel;
    else erel = (uint32_t)min(entry_abs - t_begin, (uint64_t)0x7FFFFFFFu);  // entry_abs >= t_begin
    // Stride prediction: when every record of the previous tile had the same size S
    // the starts of this tile are guessed as entry + k*S (one header check each)
    // instead of scanning for plausible headers.  Still only a guess.
    uint32_t S = st.stride;
    // A segment that has to guess its entry tries the host's expectation first (equal-sized records
    // all the way from the tail): if every header of the tile agrees, the uniform fast path below
    // takes it from the first tile on; if not, the scan guesses as before.
    const bool hinted = entry_abs == kNone && tile_k == 0 && seg.entry_hint != kNone &&
                        seg.entry_hint >= t_begin && seg.entry_hint - t_begin < (uint64_t)kTile;
    if (hinted) {
      erel = (uint32_t)(seg.entry_hint - t_begin);
      S = seg.hint_size;
    }
    bool predict = S >= 8 && erel < (uint32_t)kTile;
    const float rcpS = predict ? __frcp_rn((float)S) : 0.f;

    // ---- uniform fast path -------------------------------------------------------
    // When the previous tile held records of one size S, try the same for this tile:
    // record k would start at entry + k*S.  Every thread checks the headers of its
    // records k = tid, tid + 256, ...; if ALL of them carry size S the record chain IS
    // entry + k*S (induction from the exact entry) -- no guessing, no sub-chunks.  Any
    // other record (other size, bad SAMPLE/READ layout, timestamp regression, a counted
    // non-SAMPLE type) sends the tile to the general path below.
    bool ufp_done = false;
    if (predict && !stopped && S <= (uint32_t)kTile) {
      const uint32_t n_t = lim > erel ? (lim - erel + S - 1) / S : 0u;  // records starting in the tile
      const uint64_t in_ev = st.ev;         // interval state in front of the tile
      const uint64_t n_rows_in = st.n_rows;
      bool ok = true;
      uint32_t my_n = 0;
      for (uint32_t k = tid; k < n_t && ok; k += kThreads) {
        const uint32_t pos = erel + k * S;
        const uint64_t h = tv.at(pos);
        const uint32_t type = hdr_type(h);
        if (hdr_size(h) != S) {
          ok = false;
        } else if (type == REC_SAMPLE && h_sample) {
          if (group_layout) {
            const uint64_t nr = tv.at(pos + 16);
            if ((uint64_t)S != 40 + 8 * nr || (AGG == HBT_AGG_COUNT_DELTAS && nr != d.n_events)) ok = false;
            if (AGG == HBT_AGG_COUNT_DELTAS && ok) {
              const uint64_t ts = tv.at(pos + 8);
              if (k > 0) {
                if (tv.at(pos - S + 8) > ts) ok = false;
              } else if (ev_kind(in_ev) == EV_SAMPLE) {
                if (tv.word_abs(ev_pos(in_ev) + 8) > ts) ok = false;
              }
            }
          } else if (S != d.ip_sample_bytes) {
            ok = false;
          }
          ++my_n;
        } else if (AGG == HBT_AGG_COUNT_DELTAS || counted(type, tmask)) {
          ok = false;  // LOST / READ / context-switch records: the general path counts them
        }
      }
      if (__syncthreads_and(ok)) {
        ufp_done = true;
        my_rec += (n_t > tid) ? (n_t - tid + kThreads - 1) / kThreads : 0u;
        my_sample += my_n;
        if (AGG == HBT_AGG_IP_HISTOGRAM) {
          for (uint32_t k = tid; k < n_t; k += 2 * kThreads) {
            // two samples per thread per pass.  A list append is an atomic on the list's
            // cursor that has to return before the store can be placed: the stores are
            // deferred to the next pass / tile (pend), so nobody waits for the round trip
            pend.flush(p.hist);
            const uint32_t p0 = erel + k * S, p1 = p0 + kThreads * S;
            const bool s0 = hdr_type(tv.at(p0)) == REC_SAMPLE && h_sample;
            const bool s1 = k + kThreads < n_t && hdr_type(tv.at(p1)) == REC_SAMPLE && h_sample;
            uint32_t hs0 = 0, hs1 = 0;
            if (s0) {
              pend.ip0 = tv.at(p0 + 8);
              pend.pid0 = (uint32_t)tv.at(p0 + 16);
              pend.m0 = !hist_cache_count(cache, pend.ip0, pend.pid0, hs0);
            }
            if (s1) {
              pend.ip1 = tv.at(p1 + 8);
              pend.pid1 = (uint32_t)tv.at(p1 + 16);
              pend.m1 = !hist_cache_count(cache, pend.ip1, pend.pid1, hs1);
            }
            if (pend.m0) pend.q0 = hist_reserve(p.hist, seg.cpu, hs0, pend.l0);
            if (pend.m1) pend.q1 = hist_reserve(p.hist, seg.cpu, hs1, pend.l1);
          }
        }
        uint32_t rows_tile = 0;
        if (AGG == HBT_AGG_COUNT_DELTAS && n_t > 0) {
          // every record is a SAMPLE: record k > 0 closes the interval opened by k-1;
          // record 0 depends on what came before the tile
          const uint32_t ik = ev_kind(in_ev);
          const uint32_t first_row = ik == EV_SAMPLE ? 1u : 0u;
          rows_tile = n_t - 1 + first_row;
          const uint32_t wpr = 2 + d.n_events;
          uint64_t* out = seg.rows + (1 + n_rows_in) * (uint64_t)wpr;
          const uint64_t far_prev = ev_pos(in_ev);
          // Consecutive threads write consecutive u64s of the tile's rows.  A pass covers
          // rw_rows whole rows with the first rw_rows * wpr threads, so a thread's word
          // column rw_w (hence the field offset and the TSC transform flag) never changes
          // and its row index just advances by rw_rows: no division in the loop.
          if (rw_active) {
            const uint32_t off = rw_w < 2 ? 8u : 24u + 8u * rw_w;  // tstamp, or value[rw_w - 2] at 40 + 8 (rw_w - 2)
            const bool xf = rw_w >= 2 && ((d.tsc_transform_mask >> (rw_w - 2)) & 1);
            uint32_t cpos = erel + (rw_r0 + 1 - first_row) * S + off;  // field of the record closing row r
            const uint32_t step = rw_rows * S;
            uint64_t* o = out + rw_r0 * wpr + rw_w;
            const uint32_t ostep = rw_rows * wpr;
            for (uint32_t r = rw_r0; r < rows_tile; r += rw_rows, cpos += step, o += ostep) {
              const uint64_t cur = tv.at(cpos);
              uint64_t prv;
              if (r + 1 > first_row) prv = tv.at(cpos - S);
              else if (rowd.prev_pos[(tile_k & 1) ^ 1] == far_prev) prv = rowd.prev_words[(tile_k & 1) ^ 1][off >> 3];
              else prv = ring_word_g(ring, far_prev + off);
              uint64_t v;
              if (rw_w == 0) v = prv;
              else if (!xf) v = cur - prv;   // duration, or a plain counter delta
              else v = tsc_to_ktime(cur, d.tsc) - tsc_to_ktime(prv, d.tsc);
              *o = v;
            }
          }
          // keep the tile's last SAMPLE for the next tile's first row
          {
            const uint32_t last = erel + (n_t - 1) * S;
            if (tid < S / 8) rowd.prev_words[tile_k & 1][tid] = tv.at(last + 8 * tid);
            if (tid == 0) rowd.prev_pos[tile_k & 1] = t_begin + last;
          }
          if (tid == 0) {
            if (ik == EV_LOST) ++st.n_unknown;  // record 0 re-seeds after a LOST
            else if (ik == EV_NONE) st.first_ev = ev_make(t_begin + erel, EV_SAMPLE);  // seam
          }
        }
        if (tid == 0 && n_t > 0) {
          if (st.entry == kNone) st.entry = t_begin + erel;
          if (hinted) st.stride = S;
          st.pos = t_begin + erel + (uint64_t)n_t * S;
          if (AGG == HBT_AGG_COUNT_DELTAS) {
            const uint64_t last = t_begin + erel + (uint64_t)(n_t - 1) * S;
            st.ev = ev_make(last, EV_SAMPLE);
            st.last_sample = last;
            st.n_rows += rows_tile;
          }
        }
      }
    }

    if (hinted && !ufp_done) {  // the expectation did not hold here: guess by scanning
      erel = kNoRel;
      S = 0;
      predict = false;
    }
    const uint32_t se = min(se_full, lim);
    const bool live = !stopped && !ufp_done && sb < se;

    uint32_t my_entry, mask, smask, omask, my_x, first_ev, last_ev, my_ls, rows_local,
        unknown_local, sev, srow, rows_j, unknown_j;
    uint64_t in_ev0;
    bool owner, pending_seam, slow;
    uint32_t attempt = predict ? 0u : 1u;
    if (!ufp_done) {
#pragma unroll 1
    for (;; ++attempt) {
      // ---- phase 1: guess + walk over the own sub-chunk -------------------------
      my_entry = kNoRel;  // first record start in my sub-chunk
      owner = false;      // my_entry is the tile's exact entry
      if (live) {
        if (erel != kNoRel && erel >= sb) {
          if (erel < se) {
            my_entry = erel;
            owner = true;
          }
        } else {
          uint32_t q = sb;
          if (attempt == 0) {
            // smallest entry + k*S >= sb
            const uint32_t kk = (uint32_t)((float)(sb - erel + S - 1) * rcpS);
            q = erel + kk * S;
            if (q < sb) q += S;
            else if (q - sb >= S) q -= S;
            if (q < se && plausible(tv.at(q))) my_entry = q;
            else if (q < se) q = sb;  // prediction failed here: scan
            else q = se;              // predicted: nothing starts in my sub-chunk
          }
          if (my_entry == kNoRel && q < se) {
            // Scan: the first plausible header at or after q whose successor is plausible
            // too.  All words of the sub-chunk (it lies inside tile k) are loaded at once
            // -- independent loads, no divergent trip count -- and reduced to a bit mask
            // plus their 16-bit sizes; only the candidates' successors are looked up.
            // (the skew makes a sub-chunk kSubWords + 1 words long at most; loading a few words
            // past a short one stays inside the staging area and is masked out below)
            constexpr int kScanWords = kSubWords + 1;
            constexpr int kScanGroups = (kScanWords + 3) / 4;
            static_assert(kScanGroups <= 5, "size registers below");
            uint32_t pm = 0;
            uint64_t sz0 = 0, sz1 = 0, sz2 = 0, sz3 = 0, sz4r = 0;  // 16-bit sizes, 4 per register
#pragma unroll
            for (int w = 0; w < kScanWords; ++w) {
              const uint64_t h = lds_u64(tv.sa + sb + 8u * w);
              if (plausible(h)) pm |= 1u << w;
              const uint64_t sz = (uint64_t)hdr_size(h) << (16 * (w & 3));
              if ((w >> 2) == 0) sz0 |= sz;
              else if ((w >> 2) == 1) sz1 |= sz;
              else if ((w >> 2) == 2) sz2 |= sz;
              else if ((w >> 2) == 3) sz3 |= sz;
              else sz4r |= sz;
            }
            const uint32_t w0 = (q - sb) >> 3, w1 = (se - sb) >> 3;  // words [w0, w1)
            pm &= ((1u << w1) - 1u) & ~((1u << w0) - 1u);
            while (pm) {
              const uint32_t w = __ffs(pm) - 1;
              pm &= pm - 1;
              const uint32_t g = w >> 2;
              const uint64_t sz4 = g == 0 ? sz0 : g == 1 ? sz1 : g == 2 ? sz2 : g == 3 ? sz3 : sz4r;
              const uint32_t qq = sb + 8u * w;
              const uint32_t q2 = qq + (uint32_t)((sz4 >> (16 * (w & 3))) & 0xFFFFu);
              if (q2 < look && !plausible(tv.at(q2))) continue;
              my_entry = qq;
              break;
            }
          }
        }
      }
      mask = 0;        // bit w: a record starts at sb + 8w
      smask = 0;       // ... and it is a handled SAMPLE
      omask = 0;       // ... and it is another counted record type
      my_x = kNoRel;   // where my walk left the sub-chunk
      bool bad = false;
      uint32_t smin = kNoRel, smax = 0;
      first_ev = 0;    // tile-relative encoded events ((rel+1)<<2 | kind)
      last_ev = 0;
      my_ls = 0;       // last SAMPLE start, relative + 1
      rows_local = 0;
      unknown_local = 0;
      uint64_t prev_ts = 0;
      if (my_entry != kNoRel) {
        uint32_t pos = my_entry;
#pragma unroll 1
        while (pos < se) {
          const uint64_t h = tv.at(pos);
          const uint32_t type = hdr_type(h), size = hdr_size(h);
          if (size < 8 || (size & 7) || pos + size > 2u * kTile) {
            bad = true;
            break;
          }
          const uint32_t bit = 1u << ((pos - sb) >> 3);
          smin = min(smin, size);
          smax = max(smax, size);
          if (type == REC_SAMPLE && h_sample) {
            if (group_layout) {
              const uint64_t nr = tv.at(pos + 16);
              if ((uint64_t)size != 40 + 8 * nr ||
                  (AGG == HBT_AGG_COUNT_DELTAS && nr != d.n_events)) {
                bad = true;
                break;
              }
              if (AGG == HBT_AGG_COUNT_DELTAS) {
                const uint64_t ts = tv.at(pos + 8);
                const uint32_t lk = last_ev & 3;
                if (lk == EV_SAMPLE) {
                  if (prev_ts > ts) {
                    bad = true;
                    break;
                  }
                  ++rows_local;
                } else if (lk == EV_LOST) {
                  ++unknown_local;
                }
                last_ev = ((pos + 1) << 2) | EV_SAMPLE;
                if (first_ev == 0) first_ev = last_ev;
                my_ls = pos + 1;
                prev_ts = ts;
              }
            } else if (size != d.ip_sample_bytes) {
              bad = true;
              break;
            }
            smask |= bit;
          } else if (type == REC_READ && h_read) {
            if ((uint64_t)size != 40 + 8 * tv.at(pos + 16)) {
              bad = true;
              break;
            }
            omask |= bit;
          } else {
            if (AGG == HBT_AGG_COUNT_DELTAS && type == REC_LOST && h_lost) {
              last_ev = ((pos + 1) << 2) | EV_LOST;
              if (first_ev == 0) first_ev = last_ev;
            }
            if (counted(type, tmask)) {
              if (AGG == HBT_AGG_SWITCH_EVENTS && size < sw_min_size(type)) {
                bad = true;  // the sequential walk reports -EPERM
                break;
              }
              omask |= bit;
            }
          }
          mask |= bit;
          pos += size;
        }
        my_x = pos;
        // tell the sub-chunk the chain continues in what it should have found
        if (!bad) {
          if (pos < lim) sm.expect[sub_of_word(pos >> 3)] = pos;
          else st.tile_x[attempt] = pos;
        }
      }
      {
        const uint32_t wmin = __reduce_min_sync(0xffffffffu, smin);
        const uint32_t wmax = __reduce_max_sync(0xffffffffu, smax);
        if (lane == 0 && wmax) {
          atomicMin(&st.size_min, wmin);
          atomicMax(&st.size_max, wmax);
        }
      }
      // At a segment start without a known entry the smallest guess is the entry guess.
      if (erel == kNoRel && my_entry != kNoRel) atomicMin(&st.guess_min[attempt], my_entry);

      // DELTAS: interval state needs the last event / row count before my sub-chunk
      sev = last_ev;
      if (AGG == HBT_AGG_COUNT_DELTAS) {
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const uint32_t oe = __shfl_up_sync(0xffffffffu, sev, o);
          if (lane >= o) sev = max(sev, oe);
        }
        const uint32_t wls = __reduce_max_sync(0xffffffffu, my_ls);
        if (lane == 31) {
          sm.warp_ev[warp] = sev;
          sm.warp_ls[warp] = wls;
        }
      }
      __syncthreads();  // A: expect[], tile_x, guess_min, warp_ev are complete

      // ---- phase 2: every guess must be what the chain before it predicted ------
      if (erel == kNoRel && my_entry != kNoRel && my_entry == st.guess_min[attempt]) owner = true;
      {
        const uint32_t exp = sm.expect[tid];
        sm.expect[tid] = kNoRel;  // only I read my slot: reset it for the next round
        if (live) {
          if (owner) {
            if (exp != kNoRel && exp != my_entry) bad = true;
          } else if (exp != my_entry) {
            bad = true;  // missed a start, or guessed one the chain does not reach
          }
        } else if (exp != kNoRel && !stopped) {
          bad = true;
        }
      }
      // interval state in front of my first SAMPLE
      in_ev0 = 0;  // absolute encoding
      rows_j = rows_local;
      unknown_j = unknown_local;
      pending_seam = false;
      srow = 0;
      if (AGG == HBT_AGG_COUNT_DELTAS) {
        uint32_t pev = __shfl_up_sync(0xffffffffu, sev, 1);
        if (lane == 0) pev = 0;
        for (uint32_t w = 0; w < warp; ++w) pev = max(pev, sm.warp_ev[w]);
        in_ev0 = pev ? ev_make(t_begin + (pev >> 2) - 1, pev & 3) : st.ev;
        if ((first_ev & 3) == EV_SAMPLE) {
          const uint32_t ik = ev_kind(in_ev0);
          if (ik == EV_SAMPLE) {
            if (tv.word_abs(ev_pos(in_ev0) + 8) > tv.at((first_ev >> 2) - 1 + 8)) bad = true;
            ++rows_j;
          } else if (ik == EV_LOST) {
            ++unknown_j;
          } else {
            pending_seam = true;  // resolved by k_finalize
          }
        }
        srow = rows_j;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const uint32_t orow = __shfl_up_sync(0xffffffffu, srow, o);
          if (lane >= o) srow += orow;
        }
        if (lane == 31) sm.warp_rows[warp] = srow;
      }
      if (AGG == HBT_AGG_SWITCH_EVENTS) {
        // one proto-event per handled record: their index is a prefix sum over the threads
        rows_j = __popc(omask);
        srow = rows_j;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const uint32_t orow = __shfl_up_sync(0xffffffffu, srow, o);
          if (lane >= o) srow += orow;
        }
        if (lane == 31) sm.warp_rows[warp] = srow;
      }
      slow = __syncthreads_or(bad) != 0;  // B
      if (!slow || attempt == 1) break;
      // the stride prediction did not hold: guess again by scanning
    }
    }  // !ufp_done
    uint64_t n_rows0 = 0;
    if (ufp_done) slow = false;
    if (!stopped && !slow && !ufp_done) {
      // ---- phase 3: aggregate my records ----------------------------------------
      uint64_t row_i = 0;
      uint64_t in_ev = in_ev0;
      if (AGG == HBT_AGG_COUNT_DELTAS || AGG == HBT_AGG_SWITCH_EVENTS) {
        n_rows0 = st.n_rows;
        row_i = n_rows0 + (srow - rows_j);
        for (uint32_t w = 0; w < warp; ++w) row_i += sm.warp_rows[w];
      }
      my_rec += __popc(mask);
      my_sample += __popc(smask);
      if (AGG == HBT_AGG_IP_HISTOGRAM) {
        uint32_t m = smask;
        while (m) {
          const uint32_t pos = sb + 8u * (__ffs(m) - 1);
          m &= m - 1;
          hist_count(p.hist, cache, seg.cpu, tv.at(pos + 8), (uint32_t)tv.at(pos + 16));
        }
      }
      uint32_t m = AGG == HBT_AGG_COUNT_DELTAS ? (smask | omask) : omask;
      while (m) {
        const uint32_t pos = sb + 8u * (__ffs(m) - 1);
        m &= m - 1;
        const uint64_t h = tv.at(pos);
        const uint32_t type = hdr_type(h);
        if (AGG == HBT_AGG_COUNT_DELTAS && type == REC_SAMPLE && h_sample) {
          if (ev_kind(in_ev) == EV_SAMPLE) {
            // row = {t_prev, t - t_prev, v_i - v_prev_i}: written cooperatively below
            const uint64_t pp = ev_pos(in_ev);
            uint32_t prel = kNoRel;
            if (pp >= t_begin && pp - t_begin < 2ull * kTile) prel = (uint32_t)(pp - t_begin);
            else rowd.far_prev = pp;  // only the tile's first row can reach back this far
            rowd.d[row_i - n_rows0] = make_uint2(pos, prel);
            ++row_i;
          }
          in_ev = ev_make(t_begin + pos, EV_SAMPLE);
        } else {
          const int ci = counter_of(type, tmask);
          if (ci == HBT_CNT_LOST) {
            my_lost_sum += tv.at(pos + 16);
            ++my_lost_n;
            if (AGG == HBT_AGG_COUNT_DELTAS) in_ev = ev_make(t_begin + pos, EV_LOST);
          } else if (ci >= 0) {
            my_other += 1ull << (7 * other_field(ci));
            if (++my_other_n == 127) {  // before a 7-bit field can wrap
              flush_other(false);
              my_other_n = 0;
            }
          }
          if (AGG == HBT_AGG_SWITCH_EVENTS && ci >= 0) {
            const ulonglong2 e = sw_proto(h, d.tid_level, p.cpu_base + seg.cpu,
                                          [&](uint32_t off) { return tv.at(pos + off); });
            *reinterpret_cast<ulonglong2*>(seg.rows + (1 + row_i) * 2) = e;
            ++row_i;
          }
        }
      }
      if (AGG == HBT_AGG_COUNT_DELTAS) {
        if (unknown_j) atomicAdd(&st.n_unknown, (unsigned long long)unknown_j);
        if (pending_seam) st.first_ev = ev_make(t_begin + (first_ev >> 2) - 1, EV_SAMPLE);
      }
    }
    if ((AGG == HBT_AGG_COUNT_DELTAS || AGG == HBT_AGG_SWITCH_EVENTS) && !ufp_done) {
      __syncthreads();  // row descriptors are complete; nobody reads st.n_rows any more
      if (AGG == HBT_AGG_COUNT_DELTAS && !stopped && !slow) {
        // ---- phase 3b: write the tile's rows, consecutive threads -> consecutive words
        uint32_t trows = 0;
        for (uint32_t w = 0; w < kThreads / 32; ++w) trows += sm.warp_rows[w];
        const uint32_t wpr = 2 + d.n_events;
        uint64_t* out = seg.rows + (1 + n_rows0) * (uint64_t)wpr;
        if (rw_active) {  // same thread -> word column mapping as the uniform fast path
          const uint32_t off = rw_w < 2 ? 8u : 24u + 8u * rw_w;
          const bool xf = rw_w >= 2 && ((d.tsc_transform_mask >> (rw_w - 2)) & 1);
          uint64_t* o = out + rw_r0 * wpr + rw_w;
          for (uint32_t r = rw_r0; r < trows; r += rw_rows, o += rw_rows * wpr) {
            const uint2 rd = rowd.d[r];
            const uint64_t cur = tv.at(rd.x + off);
            uint64_t prv;
            if (rd.y != kNoRel) prv = tv.at(rd.y + off);
            else if (rowd.prev_pos[(tile_k & 1) ^ 1] == rowd.far_prev) prv = rowd.prev_words[(tile_k & 1) ^ 1][off >> 3];
            else prv = ring_word_g(ring, rowd.far_prev + off);
            uint64_t v;
            if (rw_w == 0) v = prv;
            else if (!xf) v = cur - prv;
            else v = tsc_to_ktime(cur, d.tsc) - tsc_to_ktime(prv, d.tsc);
            *o = v;
          }
        }
      }
    }
    if (tid == 0 && !ufp_done) {
      if (!stopped) {
        const uint32_t a = attempt;  // the attempt whose scatter results were verified
        if (slow) {
          // literal consume() over this tile by one thread
          ++st.slow_tiles;
          const uint32_t g = min(st.guess_min[0], st.guess_min[1]);
          if (st.pos == kNone && g != kNoRel) st.pos = t_begin + g;
          st.stride = 0;
          if (st.pos != kNone) {
            consume_sequential<AGG>(p, ring, seg, st, t_begin + lim);
            if (st.err == 0 && st.size_min == st.size_max) st.stride = st.size_min;  // exact entry + one size
          }
        } else {
          // chain bookkeeping for the next tile
          const uint32_t e = erel != kNoRel ? erel : st.guess_min[a];
          if (e != kNoRel) {
            if (st.entry == kNone && e < lim) st.entry = t_begin + e;
            if (st.tile_x[a] != kNoRel) st.pos = t_begin + st.tile_x[a];
            else if (erel == kNoRel) st.pos = t_begin + e;
          }
          if (AGG == HBT_AGG_COUNT_DELTAS) {
            uint32_t tev = 0, trows = 0, tls = 0;
            for (uint32_t w = 0; w < kThreads / 32; ++w) {
              tev = max(tev, sm.warp_ev[w]);
              tls = max(tls, sm.warp_ls[w]);
              trows += sm.warp_rows[w];
            }
            if (tev) st.ev = ev_make(t_begin + (tev >> 2) - 1, tev & 3);
            if (tls) st.last_sample = t_begin + tls - 1;
            st.n_rows += trows;
          }
          if (AGG == HBT_AGG_SWITCH_EVENTS) {
            uint32_t trows = 0;
            for (uint32_t w = 0; w < kThreads / 32; ++w) trows += sm.warp_rows[w];
            st.n_rows += trows;
          }
          st.stride = st.size_min == st.size_max ? st.size_min : 0;
        }
      }
      st.tile_x[0] = st.tile_x[1] = kNoRel;
      st.guess_min[0] = st.guess_min[1] = kNoRel;
      st.size_min = kNoRel;
      st.size_max = 0;
    }
    // every generic-proxy read of stage sa_i is ordered before the bulk copy that
    // thread 0 issues into it after this barrier
    fence_proxy_async();
    __syncthreads();  // C
    sa_i = sb_i;
  }

  // ---- segment epilogue --------------------------------------------------------
  if (AGG == HBT_AGG_IP_HISTOGRAM) {
    pend.flush(p.hist);
    for (int i = tid; i < kCacheSlots; i += kThreads) {
      const ulonglong2 e = cache.slots[i];
      if ((e.y >> 32) != 0) {
        hist_append(p.hist, seg.cpu, e.x, (uint32_t)e.y, (uint32_t)(e.y >> 32),
                    hist_hash(e.x, (uint32_t)e.y));
      }
    }
  }
  my_rec = __reduce_add_sync(0xffffffffu, my_rec);
  my_sample = __reduce_add_sync(0xffffffffu, my_sample);
  flush_other(true);
  my_lost_n = __reduce_add_sync(0xffffffffu, my_lost_n);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) my_lost_sum += __shfl_xor_sync(0xffffffffu, my_lost_sum, o);
  if (lane == 0 && my_lost_n) {
    atomicAdd(&st.counters[HBT_CNT_LOST], (unsigned long long)my_lost_sum);
    atomicAdd(&st.n_lost_recs, my_lost_n);
  }
  if (lane == 0) {
    atomicAdd(&st.n_rec, (unsigned long long)my_rec);
    atomicAdd(&st.counters[HBT_CNT_SAMPLE], (unsigned long long)my_sample);
  }
  __syncthreads();
  if (tid == 0) {
    SegResult& r = p.results[seg_index];
    r.entry = st.entry;
    r.exit = st.pos;
    r.n_records = st.n_rec;
    for (int i = 0; i < HBT_NUM_COUNTERS; ++i) r.counters[i] = st.counters[i];
    r.n_rows = st.n_rows;
    r.n_unknown = st.n_unknown;
    r.first_ev = st.first_ev;
    r.last_ev = st.ev;
    r.last_sample = st.last_sample;
    r.err = st.err;
    r.err_pos = st.err_pos;
    r.slow_tiles = st.slow_tiles;
    r.hist_overflow = 0;
    r.n_lost_recs = st.n_lost_recs;
  }
}

// Rings too small to tile (data_size < 2*kTile): one thread per CPU runs the
// literal consume() over the whole window.  Same outputs as k_decode.
template <int AGG>
__global__ void __launch_bounds__(32) k_decode_small(const DecodeParams p, uint32_t n_segs) {
  const uint32_t gt = blockIdx.x * blockDim.x + threadIdx.x;
  if (gt >= n_segs) return;
  const uint32_t g = p.run_list ? p.run_list[gt] : gt;
  const SegDesc seg = p.segs[g];
  const RingDesc ring = p.rings[seg.cpu];
  SegState st;
  st.pos = ring.tail;
  st.ev = 0;
  st.first_ev = 0;
  st.last_sample = kNone;
  st.entry = kNone;
  st.n_rec = st.n_rows = st.n_unknown = 0;
  for (int i = 0; i < HBT_NUM_COUNTERS; ++i) st.counters[i] = 0;
  st.err = 0;
  st.err_pos = 0;
  st.slow_tiles = 1;
  st.n_lost_recs = 0;
  consume_sequential<AGG>(p, ring, seg, st, ring.head);
  SegResult& r = p.results[g];
  r.entry = st.entry;
  r.exit = st.pos;
  r.n_records = st.n_rec;
  for (int i = 0; i < HBT_NUM_COUNTERS; ++i) r.counters[i] = st.counters[i];
  r.n_rows = st.n_rows;
  r.n_unknown = st.n_unknown;
  r.first_ev = st.first_ev;
  r.last_ev = st.ev;
  r.last_sample = st.last_sample;
  r.err = st.err;
  r.err_pos = st.err_pos;
  r.slow_tiles = 1;
  r.hist_overflow = 0;
  r.n_lost_recs = st.n_lost_recs;
}

// ---------------------------------------------------------------------------
// k_finalize: one thread per CPU chains the segments (exit == next entry),
// resolves the seams between segments, and produces consume()'s return value
// and the generator state after the call.
// ---------------------------------------------------------------------------
constexpr uint64_t kClampLocate = ~0ull - 1;  // clamp_head value: "locate record #max"

struct FinalizeParams {
  const RingDesc* rings;
  const SegDesc* segs;
  uint64_t* forced_out;  // per segment: true entry found for a wrongly guessed segment
  const SegResult* results;
  const CpuState* state_in;
  CpuState* state_out;
  hbt_cpu_result* out;
  uint8_t* seam_valid;   // per segment
  uint64_t* clamp_head;  // per CPU: kNone, kClampLocate, or where a re-run must stop
  int64_t* clamp_ret;    // per CPU: return value to report after such a re-run
  uint64_t* locate_from; // per CPU: record start to count from ...
  uint64_t* locate_skip; // ... and how many records (or rows) to skip to find the cut
  uint32_t* locate_mode; // per CPU: 0 = count records; 1 / 2 = count delta rows, the interval in front
                         //          of locate_from being invalid / valid
  const uint64_t* rows_budget;  // per CPU: rows the caller's sink can take (null: no limit)
  uint32_t* flags;       // [0] re-run needed
  uint8_t* cpu_done;     // per CPU: 1 = this decode's results for the CPU are final (a re-run skips it)
  uint32_t n_cpus;
  hbt_decode_desc d;
  int agg;
  uint64_t seg_limit_tile;  // kTile, or 0 for the untiled small-ring path
};

// One warp per CPU: the segment results are staged in shared memory 32 at a time by the
// whole warp (coalesced), lane 0 walks them -- the chain is sequential, but it no longer pays
// a global-memory round trip per field.
__global__ void __launch_bounds__(32) k_finalize(const FinalizeParams p) {
  const uint32_t c = blockIdx.x, lane = threadIdx.x;
  if (c >= p.n_cpus) return;
  if (p.cpu_done[c]) return;  // settled by an earlier attempt of this decode
  __shared__ SegResult s_res[32];
  __shared__ SegDesc s_seg[32];
  // Timestamps the chain walk compares, fetched by the whole warp before lane 0 walks the chunk (the walk
  // itself then touches shared memory only), and the seam rows it decides on, written by the whole warp
  // after it: global-memory latency is paid per chunk of 32 segments, not per segment.
  __shared__ uint64_t s_ts_first[32];   // tstamp of the segment's first SAMPLE (when first_ev is one)
  __shared__ uint64_t s_ts_last[32];    // tstamp of the segment's last SAMPLE (when there is one)
  __shared__ uint64_t s_seam_prev[32];  // seam row wanted: the previous SAMPLE (kPrevCarry = carry arrays); kNone = no row
  __shared__ uint64_t s_seam_pts[32];   // its tstamp
  const RingDesc ring = p.rings[c];
  const hbt_decode_desc& d = p.d;
  const CpuState& sin = p.state_in[c];
  CpuState& sout = p.state_out[c];
  hbt_cpu_result& out = p.out[c];

  uint64_t counters[HBT_NUM_COUNTERS];
  for (int i = 0; i < HBT_NUM_COUNTERS; ++i) counters[i] = sin.counters[i];
  uint64_t unknown = sin.num_unknown;
  uint64_t n_rec = 0, n_rows = 0;
  uint64_t pos = ring.tail;
  // interval state: previous sample position (kPrevCarry = carry arrays) or invalid
  bool valid = sin.carry_tstamp != kInvalidTs;
  uint64_t prev_pos = kPrevCarry;  // last SAMPLE seen (defines last_[])
  int32_t err = 0;
  bool redo = false;

  uint64_t prev_ts_val = sin.carry_tstamp;  // tstamp of the SAMPLE prev_pos names
  auto prev_ts = [&]() { return prev_ts_val; };
  auto prev_val = [&](uint32_t i) {
    return prev_pos == kPrevCarry ? sin.carry_last[i]
                                  : xform(d, i, ring_word_g(ring, prev_pos + 40 + 8ull * i));
  };

  // seam_valid: bit 0 = slot 0 of the slab holds a seam row, bit 1 = the segment's rows
  // belong to this call's output (segments behind an error do not)
  for (uint32_t s = lane; s < ring.n_seg; s += 32) p.seam_valid[ring.seg_begin + s] = 0;
  __syncwarp();
  bool stop = false;  // lane 0 left the segment loop early
  for (uint32_t s0 = 0; s0 < ring.n_seg && !stop; s0 += 32) {
    const uint32_t n_chunk = min(32u, ring.n_seg - s0);
    {
      const uint64_t* src = reinterpret_cast<const uint64_t*>(p.results + ring.seg_begin + s0);
      uint64_t* dst = reinterpret_cast<uint64_t*>(s_res);
      for (uint32_t w = lane; w < n_chunk * (uint32_t)(sizeof(SegResult) / 8); w += 32) dst[w] = src[w];
      src = reinterpret_cast<const uint64_t*>(p.segs + ring.seg_begin + s0);
      dst = reinterpret_cast<uint64_t*>(s_seg);
      for (uint32_t w = lane; w < n_chunk * (uint32_t)(sizeof(SegDesc) / 8); w += 32) dst[w] = src[w];
    }
    __syncwarp();
    if (p.agg == HBT_AGG_COUNT_DELTAS) {
      s_seam_prev[lane] = kNone;
      if (lane < n_chunk) {
        const SegResult& r = s_res[lane];
        if (ev_kind(r.first_ev) == EV_SAMPLE) s_ts_first[lane] = ring_word_g(ring, ev_pos(r.first_ev) + 8);
        if (r.last_sample != kNone) s_ts_last[lane] = ring_word_g(ring, r.last_sample + 8);
      }
      __syncwarp();
    }
    if (lane == 0) {
  for (uint32_t j = 0; j < n_chunk; ++j) {
    const uint32_t s = s0 + j;
    const uint32_t gi = ring.seg_begin + s;
    const SegResult& r = s_res[j];
    const SegDesc& sd = s_seg[j];
    const uint64_t seg_limit =
        p.seg_limit_tile ? min(sd.tile_end * p.seg_limit_tile, ring.head) : ring.head;
    if (pos >= seg_limit) {
      // the record chain jumps over this segment: nothing may start in it
      if (r.entry != kNone) {
        p.forced_out[gi] = pos;
        redo = true;
        { stop = true; break; }
      }
      continue;
    }
    if (r.entry != pos) {
      p.forced_out[gi] = pos;
      redo = true;
      { stop = true; break; }
    }
    if (n_rec + r.n_records > (uint64_t)d.max_records) {
      // consume(max_num_records) stops inside this segment: the host locates the
      // cut (record #max) and re-runs with the window clamped there
      p.clamp_head[c] = kClampLocate;
      p.clamp_ret[c] = (int64_t)d.max_records;
      p.locate_from[c] = pos;
      p.locate_skip[c] = (uint64_t)d.max_records - n_rec;
      p.locate_mode[c] = 0;
      redo = true;
      { stop = true; break; }
    }
    if (p.agg == HBT_AGG_COUNT_DELTAS && p.rows_budget && ring.override_ret != -28) {
      // Back-pressure: the generator's output ring takes rows_budget more rows; the SAMPLE whose row
      // does not fit makes handleRecordSample return -ENOSPC and consume() stop in FRONT of it
      // (PerCpuCountSampleGenerator.h:182-195, CpuEventsGroup.h:1441-1444).  When that row belongs
      // to this segment the host locates the record and re-runs with the window clamped there.
      // (a regression at the segment's seam sample is found first, below, like in the reference,
      // whose timestamp check precedes the write: the seam row is only counted when it would be written)
      uint64_t seg_rows = r.n_rows;
      if (ev_kind(r.first_ev) == EV_SAMPLE && valid &&
          !(prev_ts() > s_ts_first[j]))
        ++seg_rows;
      if (n_rows + seg_rows > p.rows_budget[c]) {
        p.clamp_head[c] = kClampLocate;
        p.clamp_ret[c] = -28;  // -ENOSPC
        p.locate_from[c] = pos;
        p.locate_skip[c] = p.rows_budget[c] - n_rows;
        p.locate_mode[c] = valid ? 2 : 1;
        redo = true;
        { stop = true; break; }
      }
    }
    // seam: the segment's first SAMPLE had no SAMPLE/LOST before it in the segment
    if (p.agg == HBT_AGG_COUNT_DELTAS && ev_kind(r.first_ev) == EV_SAMPLE) {
      const uint64_t sp = ev_pos(r.first_ev);
      const uint64_t ts = s_ts_first[j];
      if (!valid) {
        ++unknown;  // re-seed (PerCpuCountSampleGenerator.h:135-154)
      } else {
        const uint64_t pts = prev_ts();
        if (pts > ts) {
          // timestamp regression exactly at a seam: re-run up to that record
          p.clamp_head[c] = sp;
          p.clamp_ret[c] = -1;  // -EPERM
          redo = true;
          { stop = true; break; }
        }
        s_seam_prev[j] = prev_pos;  // slot 0 of the slab: written by the whole warp below
        s_seam_pts[j] = pts;
        p.seam_valid[gi] = 1;
        ++n_rows;
      }
    }
    p.seam_valid[gi] |= 2;
    n_rec += r.n_records;
    n_rows += r.n_rows;
    unknown += r.n_unknown;
    for (int i = 0; i < HBT_NUM_COUNTERS; ++i) counters[i] += r.counters[i];
    if (r.last_ev != 0) valid = ev_kind(r.last_ev) == EV_SAMPLE;
    if (r.last_sample != kNone) {
      prev_pos = r.last_sample;
      prev_ts_val = s_ts_last[j];
    }
    pos = r.exit;
    if (r.err) {
      err = r.err;
      pos = r.err_pos;
      if (p.agg == HBT_AGG_IP_HISTOGRAM && s + 1 < ring.n_seg && ring.override_ret == 0) {
        // later segments already added their samples to the shared table
        p.clamp_head[c] = r.err_pos;
        p.clamp_ret[c] = r.err;
        redo = true;
      }
      { stop = true; break; }
    }
  }
    }
    stop = __shfl_sync(0xffffffffu, stop ? 1 : 0, 0) != 0;
    __syncwarp();
    if (p.agg == HBT_AGG_COUNT_DELTAS) {
      // the seam rows of this chunk: work item = (segment, word of its row), spread over the lanes
      const uint32_t W = 2 + d.n_events;
      for (uint32_t k = lane; k < n_chunk * W; k += 32) {
        const uint32_t j = k / W, col = k - j * W;
        const uint64_t pp = s_seam_prev[j];
        if (pp == kNone) continue;
        const uint64_t sp = ev_pos(s_res[j].first_ev);
        uint64_t x;
        if (col == 0) x = s_seam_pts[j];
        else if (col == 1) x = s_ts_first[j] - s_seam_pts[j];
        else {
          const uint32_t i = col - 2;
          const uint64_t pv = pp == kPrevCarry ? sin.carry_last[i] : xform(d, i, ring_word_g(ring, pp + 40 + 8ull * i));
          x = xform(d, i, ring_word_g(ring, sp + 40 + 8ull * i)) - pv;
        }
        s_seg[j].rows[col] = x;
      }
    }
    __syncwarp();  // the staging buffers are reused by the next chunk
  }
  if (lane != 0) return;
  if (redo) {
    atomicExch(&p.flags[0], 1u);
    return;
  }

  out.n_records = n_rec;
  out.new_tail = pos;
  out.n_rows = n_rows;
  for (int i = 0; i < HBT_NUM_COUNTERS; ++i) out.counters[i] = counters[i];
  out.num_unknown_count_interval = unknown;
  if (err) {
    out.ret = err;
  } else if (ring.override_ret > 0) {
    out.ret = (int64_t)n_rec;  // stopped by max_num_records
  } else if (ring.override_ret < 0) {
    out.ret = ring.override_ret;
  } else if (pos == ring.head && n_rec < (uint64_t)d.max_records) {
    out.ret = -61;  // -ENODATA (CpuEventsGroup.h:1507-1509)
  } else {
    out.ret = (int64_t)n_rec;
  }

  // generator state after the call
  for (int i = 0; i < HBT_NUM_COUNTERS; ++i) sout.counters[i] = counters[i];
  sout.num_unknown = unknown;
  uint64_t cts = sin.carry_tstamp;
  if (p.agg == HBT_AGG_COUNT_DELTAS) {
    // count_.tstamp: tstamp of the last SAMPLE, invalid after a LOST, a regression
    // or a failed write (PerCpuCountSampleGenerator.h:122,161,192); last_[] keeps
    // the (transformed) values of the last SAMPLE whatever followed it.
    cts = valid ? prev_ts() : kInvalidTs;
    // a regression found at a seam arrives here as the override of a clamped re-run
    if (ring.override_ret == -1) cts = kInvalidTs;
    for (uint32_t i = 0; i < d.n_events; ++i) sout.carry_last[i] = prev_val(i);
    if (ring.override_ret == -28) {
      // the refused SAMPLE (it starts where the clamped window ends) had already been folded into
      // last_[] when its write failed, and the failure invalidates the interval (:165-179, :191-194)
      cts = kInvalidTs;
      for (uint32_t i = 0; i < d.n_events; ++i)
        sout.carry_last[i] = xform(d, i, ring_word_g(ring, ring.head + 40 + 8ull * i));
    }
    for (uint32_t i = d.n_events; i < HBT_MAX_EVENTS; ++i) sout.carry_last[i] = sin.carry_last[i];
  } else {
    for (uint32_t i = 0; i < HBT_MAX_EVENTS; ++i) sout.carry_last[i] = sin.carry_last[i];
  }
  sout.carry_tstamp = cts;
  out.carry_tstamp = cts;
  p.cpu_done[c] = 1;
}

// Generator state right after enable(): counters 0, last_[] = 0, count_.tstamp = seed.
__global__ void __launch_bounds__(64) k_seed_state(CpuState* st, uint32_t n_cpus, uint64_t seed) {
  const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n_cpus) return;
  CpuState s;
  for (int i = 0; i < HBT_NUM_COUNTERS; ++i) s.counters[i] = 0;
  s.num_unknown = 0;
  s.carry_tstamp = seed;
  for (int i = 0; i < HBT_MAX_EVENTS; ++i) s.carry_last[i] = 0;
  st[c] = s;
}

// Where a clamped re-run must stop (header walk by one thread per flagged CPU): the record `skip`
// records after `from` (a finite max_num_records), or the SAMPLE that would write delta row number
// `skip` + 1 counted from `from` (a full output ring).  Rows follow the generator's interval state:
// a SAMPLE closes an interval (a row) when the interval is valid, a LOST invalidates it, the next
// SAMPLE only re-seeds.
__global__ void __launch_bounds__(64)
    k_locate(const RingDesc* rings, uint32_t n_cpus, uint64_t* clamp_head, const uint64_t* locate_from,
             const uint64_t* locate_skip, const uint32_t* locate_mode, uint32_t handled_mask) {
  const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n_cpus || clamp_head[c] != kClampLocate) return;
  const RingDesc ring = rings[c];
  uint64_t pos = locate_from[c];
  if (locate_mode[c] == 0) {
    for (uint64_t k = 0; k < locate_skip[c]; ++k) pos += hdr_size(ring_word_g(ring, pos));
  } else {
    bool valid = locate_mode[c] == 2;
    uint64_t left = locate_skip[c];
    for (; pos < ring.head;) {
      const uint64_t h = ring_word_g(ring, pos);
      const uint32_t type = hdr_type(h);
      if (type == REC_SAMPLE && (handled_mask & HBT_H_SAMPLE)) {
        if (valid) {
          if (left == 0) break;  // this SAMPLE's row is the one that does not fit
          --left;
        }
        valid = true;
      } else if (type == REC_LOST && (handled_mask & HBT_H_LOST)) {
        valid = false;
      }
      pos += hdr_size(h);
    }
  }
  clamp_head[c] = pos;
}

// k_hist_aggregate: folds (cpu, bucket) lists -- distinct entries left by earlier
// decodes followed by the raw samples (count 1) and segment-cache flushes the last
// decode appended -- into distinct keys, writes them back to the front of the same list
// and sets the cursor to their number.
//
// The list is staged in shared memory; a table of 32-bit words {16-bit fingerprint,
// entry index + 1} with T >= 2n slots finds duplicates: an entry claims a slot with a
// 32-bit CAS (it becomes the key's representative) or meets a word with its fingerprint,
// compares the full key against the staged representative and, if equal, adds its weight
// to the representative's and zeroes its own.  Entries whose weight is still non-zero are
// the distinct keys.  Only native 32-bit shared-memory atomics, 4-byte probes, load <= 1/2.
#ifndef HBT_AGG_SLOTS
#define HBT_AGG_SLOTS 8192
#endif
constexpr uint32_t kAggSlots = HBT_AGG_SLOTS;   // table words of the largest size class (= 2 * list capacity)
constexpr int kAggThreads = 256;
constexpr int kAggBatch = 4;           // independent loads in flight per thread

// two staging buffers (the next list is copied in while the current one folds) + the table
__host__ __device__ constexpr size_t agg_smem_bytes(uint32_t n_hi) {
  return (size_t)2 * n_hi * sizeof(ulonglong2) + (size_t)2 * n_hi * sizeof(uint32_t);
}

__device__ __forceinline__ uint32_t agg_hash(uint64_t ip, uint32_t pid) {
  // independent of hist_hash (whose top bits chose the list)
  uint32_t x = (uint32_t)ip * 0x9E3779B1u + (uint32_t)(ip >> 32) * 0x85EBCA77u + pid * 0xC2B2AE3Du;
  x ^= x >> 16;
  x *= 0x7FEB352Du;
  x ^= x >> 15;
  x *= 0x846CA68Bu;
  x ^= x >> 16;
  return x;
}

// Lists with n_lo < n <= n_hi entries are folded by this launch (shared memory is sized
// for n_hi): the host runs the small size class first (most lists, high occupancy) and
// larger ones for the few long lists.
__global__ void __launch_bounds__(kAggThreads)
    k_hist_aggregate(const HistLists t, uint32_t n_lists, uint32_t n_lo, uint32_t n_hi,
                     const uint32_t* __restrict__ skip_if_set) {
  // queued right behind k_finalize without a host round trip: when that asked for a re-run
  // of the batch (flag set) the lists are left alone, the host restores them and decodes again
  if (skip_if_set && *skip_if_set) return;
  extern __shared__ __align__(128) uint8_t agg_smem[];
  ulonglong2* ent_buf = reinterpret_cast<ulonglong2*>(agg_smem);  // [2][n_hi]
  uint32_t* tab = reinterpret_cast<uint32_t*>(agg_smem + (size_t)2 * n_hi * sizeof(ulonglong2));
  __shared__ uint32_t s_out;
  __shared__ __align__(8) uint64_t s_bar[2];
  const uint32_t tid = threadIdx.x, lane = tid & 31;
  // This CTA owns a contiguous range of lists.  Their lengths are read 256 at a time
  // (coalesced) and the lists of this pass' size class are queued in shared memory,
  // so that skipping the others costs no dependent loads.
  __shared__ uint32_t s_q[kAggThreads], s_qlen[kAggThreads];
  __shared__ uint32_t s_qn;
  if (tid == 0) {
    mbar_init(&s_bar[0], 1);
    mbar_init(&s_bar[1], 1);
    fence_mbar_init();
  }
  const uint32_t per_cta = (n_lists + gridDim.x - 1) / gridDim.x;
  const uint32_t l_begin = blockIdx.x * per_cta;
  const uint32_t l_end = min(n_lists, l_begin + per_cta);
  uint32_t it = 0;  // lists folded so far: buffer = it & 1, mbarrier parity = (it >> 1) & 1
  for (uint32_t l0 = l_begin; l0 < l_end; l0 += kAggThreads) {
    if (tid == 0) s_qn = 0;
    __syncthreads();
    {
      const uint32_t l = l0 + tid;
      const uint32_t len = l < l_end ? min(t.cursors[l], t.cap) : 0;
      if (len > n_lo && len <= n_hi) {
        const uint32_t q = atomicAdd(&s_qn, 1u);
        s_q[q] = l;
        s_qlen[q] = len;
      }
    }
    __syncthreads();
    const uint32_t qn = s_qn;
    // one bulk async copy stages a whole list; the copy of list qi+1 runs while qi folds
    auto stage = [&](uint32_t q, uint32_t buf) {
      const uint32_t bytes = s_qlen[q] * (uint32_t)sizeof(ulonglong2);
      mbar_arrive_expect_tx(&s_bar[buf], bytes);
      bulk_g2s(ent_buf + (size_t)buf * n_hi, t.entries + (uint64_t)s_q[q] * t.cap, bytes, &s_bar[buf]);
    };
    if (tid == 0 && qn) stage(0, it & 1);
    for (uint32_t qi = 0; qi < qn; ++qi, ++it) {
      const uint32_t list = s_q[qi];
      const uint32_t n = s_qlen[qi];
      const uint32_t buf = it & 1;
      ulonglong2* ent = ent_buf + (size_t)buf * n_hi;
      if (tid == 0 && qi + 1 < qn) stage(qi + 1, buf ^ 1);
      uint32_t T = 64;  // power of two >= 2n (n <= n_hi, the table has 2 * n_hi words)
      while (T < 2 * n) T <<= 1;
      for (uint32_t i = tid * 4; i < T; i += kAggThreads * 4)
        *reinterpret_cast<uint4*>(tab + i) = make_uint4(0, 0, 0, 0);
      if (tid == 0) s_out = 0;
      mbar_wait(&s_bar[buf], (it >> 1) & 1);
      __syncthreads();
      ulonglong2* lst = t.entries + (uint64_t)list * t.cap;
      // find duplicates
      for (uint32_t i = tid; i < n; i += kAggThreads) {
        const ulonglong2 e = ent[i];
        const uint32_t pid = (uint32_t)e.y, w = (uint32_t)(e.y >> 32);
        if (w == 0) continue;  // never appended with weight 0; kept for robustness
        const uint32_t x = agg_hash(e.x, pid);
        const uint32_t word = (x & 0xFFFF0000u) | (i + 1);
        uint32_t h = x & (T - 1);
        for (;;) {
          const uint32_t old = atomicCAS(tab + h, 0u, word);
          if (old == 0) break;  // representative of its key
          if (((old ^ word) & 0xFFFF0000u) == 0) {
            const uint32_t j = (old & 0xFFFFu) - 1;
            const volatile uint64_t* rj = reinterpret_cast<const volatile uint64_t*>(ent + j);
            if (rj[0] == e.x && (uint32_t)rj[1] == pid) {
              atomicAdd(reinterpret_cast<uint32_t*>(ent + j) + 3, w);
              reinterpret_cast<uint32_t*>(ent + i)[3] = 0;  // merged: only this thread touches it
              break;
            }
          }
          h = (h + 1) & (T - 1);
        }
      }
      __syncthreads();
      // write the distinct keys back to the front of the list (one counter bump per warp)
      for (uint32_t i0 = 0; i0 < n; i0 += kAggThreads) {
        const bool in = i0 + tid < n;
        const ulonglong2 e = in ? ent[i0 + tid] : make_ulonglong2(0, 0);
        const bool used = in && (e.y >> 32) != 0;
        const uint32_t m = __ballot_sync(0xffffffffu, used);
        uint32_t wbase = 0;
        if (lane == 0 && m) wbase = atomicAdd(&s_out, (uint32_t)__popc(m));
        wbase = __shfl_sync(0xffffffffu, wbase, 0);
        if (used) lst[wbase + __popc(m & ((1u << lane) - 1))] = e;
      }
      // this buffer is refilled by a bulk copy (async proxy) after the barrier: order the
      // generic-proxy accesses above before it
      fence_proxy_async();
      __syncthreads();
      if (tid == 0) t.cursors[list] = s_out;
    }
  }
}

// Gather every list into one array of entries (cpu = cpu_base + local index).
__global__ void __launch_bounds__(128)
    k_hist_gather(const HistLists t, uint32_t n_lists, uint32_t cpu_base, hbt_hist_entry* out,
                  unsigned long long* n_out, uint64_t cap_out) {
  __shared__ unsigned long long s_base;
  for (uint32_t list = blockIdx.x; list < n_lists; list += gridDim.x) {
    const uint32_t n = min(t.cursors[list], t.cap);
    if (threadIdx.x == 0) s_base = atomicAdd(n_out, (unsigned long long)n);
    __syncthreads();
    const unsigned long long base = s_base;
    if (out) {
      const ulonglong2* lst = t.entries + (uint64_t)list * t.cap;
      for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) {
        if (base + i >= cap_out) break;
        const ulonglong2 e = lst[i];
        hbt_hist_entry h;
        h.ip = e.x;
        h.pid = (uint32_t)e.y;
        h.cpu = cpu_base + list / t.buckets;
        h.count = e.y >> 32;
        out[base + i] = h;
      }
    }
    __syncthreads();
  }
}

// The same lists as one array of the 16-byte entries they are stored as, CPU after CPU (the cpu id is
// then implied by an offset table instead of being carried by every entry: a third fewer bytes to
// bring to the host).  k_hist_offsets: a CTA per CPU sums the CPU's cursors and gives every list its
// offset inside the CPU; k_hist_cpu_scan: exclusive prefix of the per-CPU totals; k_hist_pack: copy.
__global__ void __launch_bounds__(256)
    k_hist_offsets(const HistLists t, uint32_t* __restrict__ list_off, unsigned long long* __restrict__ cpu_total) {
  __shared__ uint32_t s_warp[8];
  __shared__ uint32_t s_run;
  const uint32_t c = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_run = 0;
  __syncthreads();
  for (uint32_t b0 = 0; b0 < t.buckets; b0 += 256) {
    const uint32_t b = b0 + tid;
    const uint32_t n = b < t.buckets ? min(t.cursors[(uint64_t)c * t.buckets + b], t.cap) : 0u;
    uint32_t incl = n;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= (unsigned)o) incl += y;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t before = s_run;
    for (uint32_t w = 0; w < warp; ++w) before += s_warp[w];
    if (b < t.buckets) list_off[(uint64_t)c * t.buckets + b] = before + incl - n;
    __syncthreads();
    if (tid == 255) s_run = before + incl;
    __syncthreads();
  }
  if (tid == 0) cpu_total[c] = s_run;
}

// one CTA of 512 threads: offsets[c] = entries of the CPUs before c, offsets[n_cpus] = all (n_cpus <= 512)
__global__ void __launch_bounds__(512)
    k_hist_cpu_scan(const unsigned long long* __restrict__ cpu_total, uint32_t n_cpus,
                    unsigned long long* __restrict__ offsets) {
  __shared__ unsigned long long s[512];
  const uint32_t tid = threadIdx.x;
  s[tid] = tid < n_cpus ? cpu_total[tid] : 0ull;
  __syncthreads();
  if (tid == 0) {
    unsigned long long run = 0;
    for (uint32_t i = 0; i < n_cpus; ++i) {
      const unsigned long long v = s[i];
      offsets[i] = run;
      run += v;
    }
    offsets[n_cpus] = run;
  }
}

__global__ void __launch_bounds__(128)
    k_hist_pack(const HistLists t, uint32_t n_lists, const uint32_t* __restrict__ list_off,
                const unsigned long long* __restrict__ offsets, ulonglong2* __restrict__ out, uint64_t cap_out) {
  for (uint32_t list = blockIdx.x; list < n_lists; list += gridDim.x) {
    const uint32_t n = min(t.cursors[list], t.cap);
    const uint64_t base = offsets[list / t.buckets] + list_off[list];
    const ulonglong2* lst = t.entries + (uint64_t)list * t.cap;
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x)
      if (base + i < cap_out) out[base + i] = lst[i];
  }
}

// ---------------------------------------------------------------------------
// HBT_AGG_SWITCH_EVENTS, pass 2: proto-events -> tagstack::Event stream.
// Event count of a proto = 1 + (the previous handled record of the CPU was a LOST).
// k_sw_scan: one thread per CPU chains the segments (k_decode counted each segment's LOST
//   protos; the ones that are not the slab's last give their successor a WriteErrorsEnd)
//   from the generator state
//   (last_tstamp_, in_write_errors_), gives every segment its output offset and the
//   "previous record" of its first proto, then the CPUs' totals are prefix-summed.
// k_sw_emit: per segment, block scan over 1 + prev_is_lost and the final events.
// ---------------------------------------------------------------------------
struct SwSeg {
  uint64_t base;      // first event of the segment, relative to the CPU's first event
  uint64_t prev_ts;   // tstamp of the handled record in front of the segment (last_tstamp_)
  uint64_t last_ts;   // tstamp of the segment's last proto
  uint32_t n;         // protos of the segment that belong to this call
  uint32_t lost_inner;  // LOST protos among [0, n-1)
  uint32_t prev_lost;   // the record in front of the segment was a LOST (in_write_errors_)
  uint32_t last_lost;   // the segment's last proto is a LOST
};

struct SwParams {
  const RingDesc* rings;
  const SegDesc* segs;
  const SegResult* results;
  const uint8_t* seg_flags;  // k_finalize's seam_valid: bit 1 = segment belongs to the output
  SwSeg* sw;
  hbt_cpu_result* out;
  uint64_t* state_ts;    // per CPU last_tstamp_
  uint32_t* state_err;   // per CPU in_write_errors_
  uint64_t* offsets;     // [n_cpus + 1] first event of each CPU
  hbt_switch_event* events;
  uint32_t n_cpus, n_segs, cpu_base;
};

__device__ __forceinline__ bool sw_is_lost(uint64_t w1) { return (uint32_t)(w1 & 0xFF) == HBT_SWEV_WRITE_ERRORS_START; }

// launched as ONE CTA of 1024 threads (n_cpus <= HBT_MAX_CPUS = 512): warp w chains the segments of CPUs
// w, w + 32, ... 32 segments at a time -- every lane fetches one segment's numbers, ballots find each
// segment's non-empty predecessor and a shuffle scan gives the offsets (the first version walked a CPU's
// segments with one thread: three dependent global loads per segment, 0.2 ms with k_walk's many segments)
static __global__ void __launch_bounds__(1024) k_sw_scan(const SwParams p) {
  __shared__ uint64_t s_tot[512];
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr unsigned FULL = 0xffffffffu;
  for (uint32_t c = warp; c < 512; c += 32) {
    if (c >= p.n_cpus) {
      if (lane == 0) s_tot[c] = 0;
      continue;
    }
    const RingDesc ring = p.rings[c];
    uint64_t prev_ts = p.state_ts[c];      // of the last non-empty segment so far (warp-uniform)
    uint32_t prev_lost = p.state_err[c];
    uint64_t total = 0;
    for (uint32_t s0 = 0; s0 < ring.n_seg; s0 += 32) {
      const uint32_t s = s0 + lane;
      const bool in = s < ring.n_seg;
      const uint32_t gi = ring.seg_begin + (in ? s : 0);
      // the segment's protos that belong to this call, its last proto, and the LOST protos in
      // front of that one (each gives its successor a WriteErrorsEnd)
      uint32_t n = 0, last_lost = 0, lost_inner = 0;
      uint64_t last_ts = 0;
      if (in && (p.seg_flags[gi] & 2)) n = (uint32_t)p.results[gi].n_rows;
      if (n) {
        const uint64_t* last = p.segs[gi].rows + 2 + 2 * (uint64_t)(n - 1);  // slot 0 = unused seam slot
        last_ts = last[0];
        last_lost = sw_is_lost(last[1]);
        lost_inner = p.results[gi].n_lost_recs - last_lost;
      }
      // the non-empty segment in front of mine: the nearest lane below with protos, else the carry
      const uint32_t have = __ballot_sync(FULL, n != 0);
      const uint32_t below = have & ((1u << lane) - 1u);
      const uint32_t pl = below ? 31 - __clz(below) : 0;
      const uint64_t ts_pl = __shfl_sync(FULL, last_ts, pl);
      const uint32_t lost_pl = __shfl_sync(FULL, last_lost, pl);
      const uint64_t my_prev_ts = below ? ts_pl : prev_ts;
      const uint32_t my_prev_lost = below ? lost_pl : prev_lost;
      const uint64_t mine = n ? (uint64_t)n + lost_inner + my_prev_lost : 0;
      uint64_t incl = mine;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint64_t y = __shfl_up_sync(FULL, incl, o);
        if (lane >= (unsigned)o) incl += y;
      }
      if (in) {
        SwSeg& w = p.sw[gi];
        w.n = n;
        if (n) {
          w.last_ts = last_ts;
          w.last_lost = last_lost;
          w.lost_inner = lost_inner;
        }
        w.base = total + (incl - mine);
        w.prev_ts = my_prev_ts;
        w.prev_lost = my_prev_lost;
      }
      total += __shfl_sync(FULL, incl, 31);
      if (have) {
        const uint32_t l = 31 - __clz(have);
        prev_ts = __shfl_sync(FULL, last_ts, l);
        prev_lost = __shfl_sync(FULL, last_lost, l);
      }
    }
    if (lane == 0) {
      p.state_ts[c] = prev_ts;
      p.state_err[c] = prev_lost;
      hbt_cpu_result& o = p.out[c];
      o.n_rows = total;
      o.carry_tstamp = prev_ts;                 // last_tstamp_
      o.num_unknown_count_interval = prev_lost; // in_write_errors_
      s_tot[c] = total;
    }
  }
  __syncthreads();
  // exclusive prefix sum of the totals over the CPUs (warp 0, 16 CPUs per lane)
  if (warp == 0) {
    uint64_t loc[16], sum = 0;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      loc[k] = sum;
      sum += s_tot[lane * 16 + k];
    }
    uint64_t incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint64_t y = __shfl_up_sync(FULL, incl, o);
      if (lane >= (unsigned)o) incl += y;
    }
    const uint64_t base = incl - sum;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      const uint32_t c = lane * 16 + k;
      if (c < p.n_cpus) p.offsets[c] = base + loc[k];
    }
    if (lane == 31) p.offsets[p.n_cpus] = incl;
  }
}

static __global__ void __launch_bounds__(256) k_sw_emit(const SwParams p) {
  const uint32_t gi = blockIdx.x;
  const SwSeg w = p.sw[gi];
  if (w.n == 0) return;
  const SegDesc sd = p.segs[gi];
  const ulonglong2* slab = reinterpret_cast<const ulonglong2*>(sd.rows + 2);
  ulonglong2* out = reinterpret_cast<ulonglong2*>(p.events) + p.offsets[sd.cpu] + w.base;
  const uint64_t end_w1 = sw_word1(HBT_SWEV_WRITE_ERRORS_END, HBT_LEVEL_INVALID, p.cpu_base + sd.cpu, HBT_TAG_NA);
  __shared__ uint32_t s_warp[8];
  __shared__ uint64_t s_run;
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_run = 0;
  __syncthreads();
  for (uint32_t i0 = 0; i0 < w.n; i0 += 256) {
    const uint32_t i = i0 + tid;
    const bool in = i < w.n;
    ulonglong2 e = make_ulonglong2(0, 0);
    uint64_t prev_ts = 0;
    uint32_t prev_lost = 0;
    if (in) {
      e = slab[i];
      if (i == 0) {
        prev_ts = w.prev_ts;
        prev_lost = w.prev_lost;
      } else {
        const ulonglong2 pe = slab[i - 1];
        prev_ts = pe.x;
        prev_lost = sw_is_lost(pe.y);
      }
    }
    const uint32_t cnt = in ? 1u + prev_lost : 0u;
    uint32_t inc = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += v;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t before = inc - cnt;
    for (uint32_t k = 0; k < warp; ++k) before += s_warp[k];
    uint64_t pos = s_run + before;
    if (in) {
      if (prev_lost) out[pos++] = make_ulonglong2(e.x, end_w1);       // pendingRecordLost_: End(tstamp)
      if (sw_is_lost(e.y)) e.x = (prev_lost ? e.x : prev_ts) + 1;    // Start(last_tstamp_ + 1)
      out[pos] = e;
    }
    __syncthreads();
    if (tid == 255) s_run += before + cnt;  // thread 255 holds the chunk's total
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------
// k_sw_merge: one level of the binary Combinator tree (hbt/src/tagstack/Stream.h:114-248)
// over the per-CPU event streams: runs of `run` CPUs are merged pairwise, (0,1), (2,3), ...
// like makeBinaryTree pairs its nodes; an odd run at the end is copied.  The order is
// Event::happenedBefore (Event.h:205-213): timestamp first, then comp_unit_id.  Every
// thread produces kMergeK consecutive outputs: a merge-path binary search finds where its
// diagonal cuts the two inputs, then it merges sequentially.  Exactly the Combinator's
// output when each CPU's stream is timestamp-ordered, which the Combinator itself assumes
// (its DCHECKs, Stream.h:139-141).
// ---------------------------------------------------------------------------
constexpr uint32_t kMergeK = 8;                 // outputs per thread
constexpr uint32_t kMergeTile = 256 * kMergeK;  // outputs per CTA

__device__ __forceinline__ bool ev_before(const ulonglong2& a, const ulonglong2& b) {
  if (a.x != b.x) return a.x < b.x;
  return (uint32_t)((a.y >> 16) & 0xFFFF) < (uint32_t)((b.y >> 16) & 0xFFFF);
}

// merge path: how many of the first d outputs of merge(A[0..la), B[0..lb)) come from A
template <class TA, class TB>
__device__ __forceinline__ uint64_t merge_path(TA A, uint64_t la, TB B, uint64_t lb, uint64_t d) {
  uint64_t il = d > lb ? d - lb : 0, ih = min(d, la);
  while (il < ih) {
    const uint64_t mid = (il + ih) >> 1;
    if (ev_before(A[mid], B[d - 1 - mid])) il = mid + 1;
    else ih = mid;
  }
  return il;
}

// tile_prefix[p] = first CTA of pair p (host-computed from the offsets), [n_pairs] = grid size.
// A CTA merges kMergeTile consecutive outputs of its pair.  Where the tile's diagonals cut the pair's
// two inputs is found by k_sw_cuts beforehand, one thread per tile: the ~20 dependent probes of that
// binary search in global memory used to be most of a merge CTA's lifetime, with 254 of its threads
// waiting at a barrier.
struct MergePair {
  uint64_t a0, la, lb;   // first input event of the pair, lengths of its two runs
  uint32_t first_tile;   // first CTA of the pair
  uint32_t last;         // 1: the CTA asked for is the pair's last
};
__device__ __forceinline__ MergePair merge_pair_of(uint32_t tile, const uint64_t* __restrict__ offsets, uint32_t n_cpus,
                                                   uint32_t run, const uint32_t* __restrict__ tile_prefix,
                                                   uint32_t n_pairs) {
  uint32_t lo = 0, hi = n_pairs - 1;
  while (lo < hi) {  // largest p with tile_prefix[p] <= tile
    const uint32_t mid = (lo + hi + 1) >> 1;
    if (tile_prefix[mid] <= tile) lo = mid;
    else hi = mid - 1;
  }
  MergePair m;
  m.a0 = offsets[min(2 * lo * run, n_cpus)];
  const uint64_t b0 = offsets[min((2 * lo + 1) * run, n_cpus)];
  const uint64_t b1 = offsets[min((2 * lo + 2) * run, n_cpus)];
  m.la = b0 - m.a0;
  m.lb = b1 - b0;
  m.first_tile = tile_prefix[lo];
  m.last = tile + 1 == tile_prefix[lo + 1];
  return m;
}

static __global__ void __launch_bounds__(256)
    k_sw_cuts(const ulonglong2* __restrict__ in, const uint64_t* __restrict__ offsets, uint32_t n_cpus, uint32_t run,
              const uint32_t* __restrict__ tile_prefix, uint32_t n_pairs, uint32_t n_tiles,
              uint64_t* __restrict__ cuts) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_tiles) return;
  const MergePair m = merge_pair_of(t, offsets, n_cpus, run, tile_prefix, n_pairs);
  const ulonglong2* A = in + m.a0;
  cuts[t] = merge_path(A, m.la, A + m.la, m.lb, (uint64_t)(t - m.first_tile) * kMergeTile);
}

// The two input pieces of the tile (together exactly one tile) are staged in shared memory with
// coalesced loads, every thread merges its kMergeK outputs from there into registers, and the outputs
// leave through the same shared memory (padded: one slot after every kMergeK, so that the threads'
// strided stores spread over the banks) as coalesced stores.
static __global__ void __launch_bounds__(256)
    k_sw_merge(const ulonglong2* __restrict__ in, ulonglong2* __restrict__ out,
               const uint64_t* __restrict__ offsets, uint32_t n_cpus, uint32_t run,
               const uint32_t* __restrict__ tile_prefix, uint32_t n_pairs, const uint64_t* __restrict__ cuts) {
  __shared__ ulonglong2 s_ev[kMergeTile + kMergeTile / kMergeK];
  const uint32_t tid = threadIdx.x;
  const MergePair m = merge_pair_of(blockIdx.x, offsets, n_cpus, run, tile_prefix, n_pairs);
  const ulonglong2* A = in + m.a0;
  const ulonglong2* B = A + m.la;
  const uint64_t d0 = (uint64_t)(blockIdx.x - m.first_tile) * kMergeTile;
  const uint64_t d1 = min(d0 + kMergeTile, m.la + m.lb);
  const uint64_t i0 = cuts[blockIdx.x], i1 = m.last ? m.la : cuts[blockIdx.x + 1];
  const uint32_t na = (uint32_t)(i1 - i0), n = (uint32_t)(d1 - d0), nb = n - na;
  // element k of the tile lives at k + k / kMergeK: the threads walk the two runs at strides of a few
  // elements, which without the padding puts a quarter-warp's 16-byte loads on the same banks
  auto P = [](uint32_t k) { return k + k / kMergeK; };
  for (uint32_t i = tid; i < n; i += 256) s_ev[P(i)] = i < na ? A[i0 + i] : B[(d0 - i0) + (i - na)];
  __syncthreads();
  auto sA = [&](uint32_t i) -> const ulonglong2& { return s_ev[P(i)]; };
  auto sB = [&](uint32_t i) -> const ulonglong2& { return s_ev[P(na + i)]; };
  const uint32_t dd = min(tid * kMergeK, n);
  uint32_t ai, bi;
  {  // merge path inside the tile
    uint32_t il = dd > nb ? dd - nb : 0, ih = min(dd, na);
    while (il < ih) {
      const uint32_t mid = (il + ih) >> 1;
      if (ev_before(sA(mid), sB(dd - 1 - mid))) il = mid + 1;
      else ih = mid;
    }
    ai = il;
    bi = dd - il;
  }
  const uint32_t cnt = min(kMergeK, n - dd);
  ulonglong2 o[kMergeK];
  // the heads of the two runs stay in registers: one shared-memory load per output
  ulonglong2 ha = ai < na ? sA(ai) : make_ulonglong2(~0ull, ~0ull);
  ulonglong2 hb = bi < nb ? sB(bi) : make_ulonglong2(~0ull, ~0ull);
#pragma unroll
  for (uint32_t j = 0; j < kMergeK; ++j) {
    if (j < cnt) {
      const bool take_a = bi >= nb || (ai < na && ev_before(ha, hb));
      o[j] = take_a ? ha : hb;
      if (take_a) {
        ++ai;
        if (ai < na) ha = sA(ai);
      } else {
        ++bi;
        if (bi < nb) hb = sB(bi);
      }
    }
  }
  __syncthreads();  // every thread has read its inputs
#pragma unroll
  for (uint32_t j = 0; j < kMergeK; ++j)
    if (j < cnt) s_ev[dd + j + tid] = o[j];  // P(dd + j)
  __syncthreads();
  ulonglong2* dst = out + m.a0 + d0;
  for (uint32_t i = tid; i < n; i += 256) dst[i] = s_ev[P(i)];
}

}  // namespace hbt
