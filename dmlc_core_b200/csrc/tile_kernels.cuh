/*
 * ParseTiles: the one-launch form of the text -> CSR parse (sm_100a).
 *
 * The two-pass walk of text_kernels.cuh reads the chunk twice and runs its whole front-end
 * (classify, compact, pair tokens) twice: once to count, once to emit.  Here every byte is
 * classified once, the way the reference makes one pass over a chunk
 * (TextParserBase::FillData, reference src/data/text_parser.h:93-134):
 *
 *   stage    a CTA takes the next tile of the chunk off a ticket counter and brings its text
 *            (16 bytes of left halo + tile + overhang) into shared memory with one bulk
 *            asynchronous copy (cp.async.bulk, completion on an mbarrier; this GPU's memory only)
 *   collect  each of the CTA's 8 warps walks its sub-range of the tile out of shared memory,
 *            pairs tokens into roles (the same TokenRoles rules as the two-pass walk) and files
 *            token positions, as 16-bit offsets into the staged window, into per-role queues
 *            that are large enough for the whole sub-range; this also yields the five counts
 *   place    the counts of the 8 warps are scanned inside the CTA and the tile's totals go
 *            through a decoupled look-back over the tiles in front of it (status words in
 *            global memory: aggregate first, inclusive prefix once known), which tells every
 *            warp where its rows / entries start in the CSR arrays
 *   convert  the queues are drained 32 tokens at a time, one role at a time: every lane runs
 *            the same converter on text that is still in shared memory, every store is 32
 *            consecutive elements of one array
 *
 * LibSVM / LibFM: a warp owns the LINES that start in its sub-range (they may run on into the
 * overhang).  CSV (label column 0 or none): rows need no pairing, so a warp owns the FIELDS
 * that start in its sub-range; how many values the row open at the tile's start already holds
 * travels through the look-back as well, which is all the column ids need.
 *
 * What does not fit -- a line that runs past the staged overhang, a sub-range with more tokens
 * or rows than its queues hold, CSV with a label column > 0 -- raises kErrFallback (or is never
 * sent here); the host then parses that chunk with the two-pass kernels.  Nothing is lost, it is
 * only slower.
 */
#ifndef DMLC_CORE_B200_TILE_KERNELS_CUH_
#define DMLC_CORE_B200_TILE_KERNELS_CUH_

#include "text_kernels.cuh"

namespace b200p {

constexpr int kTileWarps = 8;
constexpr int kTileThreads = kTileWarps * 32;
constexpr uint32_t kTileHalo = 16;           /* bytes staged in front of the tile */
constexpr uint32_t kStatusWords = 3;         /* 64-bit words per tile and kind (aggregate / prefix) */
constexpr uint32_t kTileHeaderBytes = 256;   /* ticket counter in front of the status words */

struct TileParams {
  const uint8_t *text;       /* 16-byte aligned */
  uint32_t size;
  uint32_t sub_bytes;        /* bytes of text per warp, multiple of kStepBytes */
  uint32_t over_bytes;       /* staged behind the tile, multiple of kStepBytes */
  uint32_t num_tiles;
  int32_t label_column;      /* CSV: 0 or negative */
  uint32_t four, sixteen;    /* see WalkParams */
  uint32_t cap_pool;         /* per-warp queue capacities, in entries */
  uint32_t cap_rows;
  unsigned char *scratch;    /* kTileHeaderBytes + num_tiles * 2 * kStatusWords * 8 bytes, zeroed */
  ChunkSummary *summary;     /* zeroed */
  unsigned long long *offset;
  float *label;
  float *weight;
  unsigned long long *qid;
  void *field;
  void *index;
  float *value;
  unsigned long long out_rows;   /* capacities of the arrays, in elements */
  unsigned long long out_nnz;
};

/* shared-memory bytes one CTA needs; mirrored by the carve-up at the top of ParseTiles */
__host__ __device__ inline uint32_t TileWindowBytes(const TileParams &p) {
  return kTileHalo + kTileWarps * p.sub_bytes + p.over_bytes;
}
template <int FMT>
__host__ __device__ inline uint32_t TileWarpQueueBytes(const TileParams &p) {
  uint32_t b = p.cap_pool * 2u + 2u * p.cap_rows * 2u;                 /* pool, row position, row nnz */
  if (FMT != kCSV) b += kTokenQueue * 4u + p.cap_rows * 2u;             /* token ring, weight */
  if (FMT == kLibSVM) b += p.cap_rows * 2u;                             /* qid */
  if (FMT == kLibFM) b += (p.cap_pool / 2u) * 2u;                       /* field */
  return (b + 15u) & ~15u;
}

struct TileShared;
template <int FMT>
__host__ __device__ inline uint32_t TileSmemBytes(const TileParams &p);

struct TileShared {
  uint8_t lut[256];
  unsigned long long bar;                       /* mbarrier of the bulk copy */
  uint32_t tile;
  uint32_t flags;
  uint32_t count[kTileWarps][8];                /* per warp: 5 counters, 5 = tail, 6 = has_row */
  uint32_t base[kTileWarps][8];                 /* per warp: 5 global bases, 5 = values in the open row */
};

template <int FMT>
__host__ __device__ inline uint32_t TileSmemBytes(const TileParams &p) {
  return ((static_cast<uint32_t>(sizeof(TileShared)) + 127u) & ~127u) + ((TileWindowBytes(p) + 15u) & ~15u) +
         kTileWarps * TileWarpQueueBytes<FMT>(p);
}

/* ---------------------------------------------------------------- text served from the tile -- */
struct TileText {
  const uint8_t *win;     /* shared memory; win[0] is the byte at absolute position win_lo */
  uint32_t win_lo;        /* multiple of 16 */
  uint32_t win_hi;        /* every byte of [win_lo, win_hi) is in `win` (NUL past the chunk) */
  Text global;            /* the whole chunk, for what the window cannot serve */
  __device__ __forceinline__ uint32_t byte(uint32_t i) const { return win[i - win_lo]; }
};

/* The same tile addressed by offsets into the window (what the role queues hold): the text
   source the converters of the convert phase run on.  Windows come out of shared memory; the
   byte-by-byte general converters, and windows that would run past the staged bytes, read the
   chunk in global memory. */
struct TileRelText {
  const uint8_t *win;
  uint32_t win_bytes;     /* staged bytes */
  uint32_t win_lo;
  Text global;
  __device__ __forceinline__ uint32_t at(uint32_t rel) const { return global.at(win_lo + rel); }
  __device__ __forceinline__ bool Window16(uint32_t rel, Window *w) const {
    const uint32_t a = rel & ~3u;
    if (a + 20u > win_bytes) return global.Window16(win_lo + rel, w);
    const uint32_t *src = reinterpret_cast<const uint32_t *>(win + a);
    const uint32_t w0 = src[0], w1 = src[1], w2 = src[2], w3 = src[3], w4 = src[4];
    const uint32_t k = rel & 3u;
    w->x0 = BytesFrom(w0, w1, k);
    w->x1 = BytesFrom(w1, w2, k);
    w->x2 = BytesFrom(w2, w3, k);
    w->x3 = BytesFrom(w3, w4, k);
    return true;
  }
  __device__ __forceinline__ bool Tail8(uint32_t rel, uint32_t *x4, uint32_t *x5) const {
    const uint32_t a = rel & ~3u, k = rel & 3u;
    if (a + 28u > win_bytes) return global.Tail8(win_lo + rel, x4, x5);
    const uint32_t *src = reinterpret_cast<const uint32_t *>(win + a);
    const uint32_t w4 = src[4], w5 = src[5], w6 = src[6];
    *x4 = BytesFrom(w4, w5, k);
    *x5 = BytesFrom(w5, w6, k);
    return true;
  }
};

/* ---------------------------------------------------------------- mbarrier / bulk copy ------- */
__device__ __forceinline__ uint32_t SharedAddr(const void *p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void BarrierInit(unsigned long long *bar, uint32_t arrivals) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(SharedAddr(bar)), "r"(arrivals));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void BarrierExpect(unsigned long long *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(SharedAddr(bar)), "r"(bytes)
               : "memory");
}
/* global -> shared, `bytes` a multiple of 16, both addresses 16-byte aligned; completes on `bar` */
__device__ __forceinline__ void BulkLoad(void *dst, const void *src, uint32_t bytes,
                                         unsigned long long *bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          SharedAddr(dst)),
      "l"(src), "r"(bytes), "r"(SharedAddr(bar))
      : "memory");
}
__device__ __forceinline__ void BarrierWait(unsigned long long *bar, uint32_t phase) {
  uint32_t done = 0;
  while (done == 0u) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(SharedAddr(bar)), "r"(phase)
        : "memory");
  }
}

/* ---------------------------------------------------------------- look-back status words ----- */
/* word 0: flag | rows | nnz     word 1: flag | values | weights     word 2: flag | stop | qids | tail
   flag: bit 63 = written.  All counts are below 2^30 (a chunk is at most 1 GiB). */
constexpr unsigned long long kFlagBit = 1ull << 63;
constexpr unsigned long long kStopBit = 1ull << 62;
__device__ __forceinline__ unsigned long long Pack2(uint32_t hi, uint32_t lo) {
  return kFlagBit | (static_cast<unsigned long long>(hi) << 31) | lo;
}
__device__ __forceinline__ uint32_t PackedHi(unsigned long long w) {
  return static_cast<uint32_t>(w >> 31) & 0x7fffffffu;
}
__device__ __forceinline__ uint32_t PackedLo(unsigned long long w) {
  return static_cast<uint32_t>(w) & 0x7fffffffu;
}
/* Every status word carries its own "written" flag and is stored with one 64-bit store, so
   readers need no ordering between the words: a slot counts as published once all of its words
   show the flag.  Loads and stores are relaxed at GPU scope (served by L2, never by a stale L1
   line); no fence is involved on either side. */
__device__ __forceinline__ unsigned long long LoadStatus(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void StoreStatus(unsigned long long *p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

struct TileTotals {
  uint32_t c[kNumCounters];
  uint32_t tail;      /* aggregate: values after the tile's last row start (all of them if it has
                         none); prefix: values the row open at the tile's end holds */
  bool stop;          /* aggregate: the tile has a row start; prefix: always */
};

__device__ __forceinline__ void PublishStatus(unsigned long long *slot, const TileTotals &t) {
  StoreStatus(slot + 0, Pack2(t.c[kRows], t.c[kNnz]));
  StoreStatus(slot + 1, Pack2(t.c[kValues], t.c[kWeights]));
  StoreStatus(slot + 2, Pack2(t.c[kQids], t.tail) | (t.stop ? kStopBit : 0ull));
}

/*
 * Exclusive prefix of tile `tile` over the tiles in front of it, by warp 0 of the CTA (all 32
 * lanes call).  Lane i looks at tile (first - i): its inclusive prefix if published, else its
 * aggregate; the walk ends at the nearest prefix.  Tiles are numbered by a ticket taken when
 * the CTA starts, so every tile waited for is running or finished -- the wait cannot deadlock.
 * A bounded number of polls guards the GPU against a logic error all the same.
 */
__device__ __forceinline__ bool LookBack(const TileParams &prm, uint32_t tile, int lane,
                                         TileTotals *excl) {
  const unsigned long long *agg = reinterpret_cast<const unsigned long long *>(prm.scratch + kTileHeaderBytes);
  const unsigned long long *pre = agg + static_cast<size_t>(prm.num_tiles) * kStatusWords;
  uint32_t sum[kNumCounters] = {0u, 0u, 0u, 0u, 0u};
  uint32_t tail = 0;
  bool tail_open = true;     /* no row start met yet on the way back */
  int first = static_cast<int>(tile) - 1;
  uint32_t polls = 0;
  while (first >= 0) {
    const int j = first - lane;
    unsigned long long w0 = kFlagBit, w1 = kFlagBit, w2 = kFlagBit | kStopBit;
    int state = 2;           /* 1 = aggregate, 2 = prefix; in front of tile 0: prefix zero */
    if (j >= 0) {
      /* all six words are requested before any is looked at: one trip to L2 per poll */
      const unsigned long long *ps = pre + static_cast<size_t>(j) * kStatusWords;
      const unsigned long long *as = agg + static_cast<size_t>(j) * kStatusWords;
      const unsigned long long p0 = LoadStatus(ps), p1 = LoadStatus(ps + 1), p2 = LoadStatus(ps + 2);
      const unsigned long long a0 = LoadStatus(as), a1 = LoadStatus(as + 1), a2 = LoadStatus(as + 2);
      if ((p0 & p1 & p2 & kFlagBit) != 0ull) {
        w0 = p0; w1 = p1; w2 = p2;
      } else if ((a0 & a1 & a2 & kFlagBit) != 0ull) {
        state = 1;
        w0 = a0; w1 = a1; w2 = a2;
      } else {
        state = 0;
      }
    }
    const uint32_t ready_m = __ballot_sync(0xffffffffu, state != 0);
    const uint32_t prefix_m = __ballot_sync(0xffffffffu, state == 2);
    const int stop_lane = prefix_m != 0u ? __ffs(prefix_m) - 1 : 31;
    const uint32_t need_m = stop_lane == 31 ? 0xffffffffu : ((2u << stop_lane) - 1u);
    if ((ready_m & need_m) != need_m) {
      if (++polls > (1u << 22)) return false;
      __nanosleep(100);
      continue;
    }
    const bool counted = lane <= stop_lane;
    const bool my_stop = counted && (state == 2 || (w2 & kStopBit) != 0ull);
    sum[kRows] += __reduce_add_sync(0xffffffffu, counted ? PackedHi(w0) : 0u);
    sum[kNnz] += __reduce_add_sync(0xffffffffu, counted ? PackedLo(w0) : 0u);
    sum[kValues] += __reduce_add_sync(0xffffffffu, counted ? PackedHi(w1) : 0u);
    sum[kWeights] += __reduce_add_sync(0xffffffffu, counted ? PackedLo(w1) : 0u);
    sum[kQids] += __reduce_add_sync(0xffffffffu, counted ? (PackedHi(w2) & 0x3fffffffu) : 0u);
    /* values of the open row: tails of the nearest tiles up to and including the first that
       holds a row start (or is a prefix) */
    const uint32_t stops = __ballot_sync(0xffffffffu, my_stop);
    const int tail_lane = stops != 0u ? __ffs(stops) - 1 : 31;
    tail += __reduce_add_sync(0xffffffffu, (tail_open && counted && lane <= tail_lane) ? PackedLo(w2) : 0u);
    if (stops != 0u) tail_open = false;
    if (prefix_m != 0u) break;
    first -= 32;
  }
#pragma unroll
  for (int c = 0; c < kNumCounters; ++c) excl->c[c] = sum[c];
  excl->tail = tail;
  excl->stop = true;
  return true;
}

/* first end-of-line byte in [from, to) of the staged window: its position or 0xffffffff */
__device__ __forceinline__ uint32_t FindEolTile(const TileText &tt, uint32_t from, uint32_t to, int lane) {
  for (uint32_t base = from & ~uint32_t(kStepBytes - 1); base < to; base += kStepBytes) {
    const uint32_t first = base + 16u * lane;
    uint32_t e = 0u;
    if (first >= tt.win_lo && first + 16u <= tt.win_hi) {   /* the step holding `from` may begin in front of the window */
      const uint4 v = *reinterpret_cast<const uint4 *>(tt.win + (first - tt.win_lo));
      e = Gather4(WordEol(v.x)) | (Gather4(WordEol(v.y)) << 4) | (Gather4(WordEol(v.z)) << 8) |
          (Gather4(WordEol(v.w)) << 12);
    }
    e &= KeepMask(first, from, to);
    const uint32_t who = __ballot_sync(0xffffffffu, e != 0u);
    if (who != 0u) {
      const int src = __ffs(who) - 1;
      const uint32_t es = __shfl_sync(0xffffffffu, e, src);
      return base + 16u * src + (__ffs(es) - 1);
    }
  }
  return 0xffffffffu;
}

/* per-warp queues, carved from dynamic shared memory */
template <int FMT>
struct TileQueues {
  uint32_t *tokens;     /* file-order ring of the front-end (LibSVM / LibFM) */
  uint16_t *pool;       /* index positions from the front, value positions from the back */
  uint16_t *field;      /* LibFM: field position of index entry i */
  uint16_t *row_at;     /* label token of row i (CSV: first field of the row) */
  uint16_t *row_nnz;    /* entries of this warp in front of row i */
  uint16_t *weight;
  uint16_t *qid;
};

template <int FMT>
__device__ __forceinline__ TileQueues<FMT> CarveQueues(uint8_t *at, const TileParams &prm) {
  TileQueues<FMT> q;
  q.tokens = reinterpret_cast<uint32_t *>(at);
  if (FMT != kCSV) at += kTokenQueue * 4u;
  q.pool = reinterpret_cast<uint16_t *>(at);
  at += prm.cap_pool * 2u;
  q.row_at = reinterpret_cast<uint16_t *>(at);
  at += prm.cap_rows * 2u;
  q.row_nnz = reinterpret_cast<uint16_t *>(at);
  at += prm.cap_rows * 2u;
  q.weight = reinterpret_cast<uint16_t *>(at);
  if (FMT != kCSV) at += prm.cap_rows * 2u;
  q.qid = reinterpret_cast<uint16_t *>(at);
  if (FMT == kLibSVM) at += prm.cap_rows * 2u;
  q.field = reinterpret_cast<uint16_t *>(at);
  return q;
}

/* what a warp knows after collecting its sub-range */
struct Collected {
  uint32_t n_rows, n_index, n_value, n_weight, n_qid;
  uint32_t tail;       /* CSV: values after the last row start (all values when there is none) */
  bool has_row;
  bool overflow;       /* the sub-range does not fit this kernel: the chunk goes to the two-pass path */
  bool dense;          /* ... because a queue ran full */
  uint32_t error;      /* lane-local error bits */
};

#ifdef B200P_CHECKED
#define B200P_TILE_CHECK(errvar, cond) \
  if (!(cond)) { (errvar) |= kErrInternal; } else
#else
#define B200P_TILE_CHECK(errvar, cond)
#endif

/* ---------------------------------------------------------------- collect: LibSVM / LibFM ---- */
template <int FMT, bool QID>
__device__ __forceinline__ void CollectTokens(const TileParams &prm, const TileText &tt,
                                              const uint8_t *lut, const TileQueues<FMT> &q,
                                              uint32_t lo, uint32_t hi, int lane, Collected *out) {
  Collected c;
  c.n_rows = c.n_index = c.n_value = c.n_weight = c.n_qid = 0u;
  c.tail = 0u;
  c.has_row = false;
  c.overflow = false;
  c.dense = false;
  c.error = 0u;
  *out = c;
  if (lo >= hi) return;

  /* owned bytes [begin, end): begin = first line start at or after lo, end = first line start
     at or after hi -- the cut FillData makes per thread (text_parser.h:104-118), looking forward */
  uint32_t begin = 0;
  if (lo != 0u) {
    const uint32_t e = FindEolTile(tt, lo - 1u, hi - 1u, lane);
    begin = (e == 0xffffffffu) ? hi : e + 1u;
  }
  if (begin >= hi) return;

  const uint32_t size = tt.global.size;
  uint32_t fill = 0;
  uint32_t end = size;
  uint32_t carry = 0x20u;        /* class nibbles of the two bytes in front: an EOL before `begin` */
  int carry_pos = 0;
  bool carry_first = false;
  uint32_t carry_at = 0;
  bool done = false;
  const uint32_t lt = LaneMaskLt();
  const uint32_t me = 1u << lane;
  const uint8_t *const my_win = tt.win + 16u * lane - tt.win_lo;   /* my_win + base = my 16 bytes */
  /* the role queues as offsets (in entries) from the pool: they are carved from one block */
  uint16_t *const q16 = q.pool;
  const uint32_t o_pool = 0u;
  const uint32_t o_row_at = static_cast<uint32_t>(q.row_at - q.pool), o_row_nnz = static_cast<uint32_t>(q.row_nnz - q.pool);
  const uint32_t o_weight = static_cast<uint32_t>(q.weight - q.pool), o_qid = static_cast<uint32_t>(q.qid - q.pool);
  const uint32_t o_field = static_cast<uint32_t>(q.field - q.pool);
  const uint32_t room_pool = prm.cap_pool - 64u, room_rows = prm.cap_rows - 32u, room_field = prm.cap_pool / 2u - 32u;

  for (uint32_t base = begin & ~uint32_t(kStepBytes - 1); !done; base += kStepBytes) {
    if (base + kStepBytes > tt.win_hi) {
      /* the line runs past what was staged.  At the end of the chunk that is simply the end
         (win_hi covers it, NUL-padded); anywhere else this kernel gives up on the chunk. */
      c.overflow = true;
      break;
    }
    const uint32_t first_byte = base + 16u * lane;
    const bool interior = base >= begin && base + kStepBytes <= hi - 1u;
    const uint4 now = *reinterpret_cast<const uint4 *>(my_win + base);
    uint32_t m0, m1;
    uint32_t k0 = 0xffffffffu, k1 = 0xffffffffu;
    ClassifyLut4(now, lut, prm.sixteen, &m0, &m1);
    if (!interior) {
      const uint32_t cm = KeepMask(first_byte, hi - 1u, size);
      const uint32_t c0 = (m0 >> 1) & kNibD & SpreadNibbles8(cm & 0xffu);
      const uint32_t c1 = (m1 >> 1) & kNibD & SpreadNibbles8(cm >> 8);
      uint32_t off = 0xffffffffu;
      if (c0 != 0u) {
        off = (__ffs(c0) - 1) >> 2;
      } else if (c1 != 0u) {
        off = 8u + ((__ffs(c1) - 1) >> 2);
      }
      const uint32_t who = __ballot_sync(0xffffffffu, off != 0xffffffffu);
      if (who != 0u) {
        const int src = __ffs(who) - 1;
        end = base + 16u * src + __shfl_sync(0xffffffffu, off, src) + 1u;
        done = true;
      } else if (base + kStepBytes >= size) {
        done = true;
      }
      const uint32_t km = KeepMask(first_byte, begin, end);
      k0 = SpreadNibbles8(km & 0xffu);
      k1 = SpreadNibbles8(km >> 8);
    }
    uint32_t up = __shfl_up_sync(0xffffffffu, m1 >> 24, 1);
    if (lane == 0) up = carry;
    carry = __shfl_sync(0xffffffffu, m1 >> 24, 31);
    const uint32_t d0 = m0 & kNibD & k0, d1 = m1 & kNibD & k1;
    const uint32_t s0 = d0 & ~((d0 << 4) | ((up >> 4) & 1u));
    const uint32_t s1 = d1 & ~((d1 << 4) | (d0 >> 28));
    const uint32_t mine = __popc(s0) + __popc(s1);
    const uint32_t incl = WarpInclusiveSum(mine, lane);
    uint32_t slot = fill + incl - mine;
    const uint32_t p_lo = up | (m0 << 8), p_mid = (m0 >> 24) | (m1 << 8), p_hi = m1 >> 24;
    const uint32_t first4 = 4u * first_byte;
    uint32_t bits = s0;
    while (bits != 0u) {
      const uint32_t b = __ffs(bits) - 1;
      bits &= bits - 1u;
      B200P_TILE_CHECK(c.error, slot < kTokenQueue)
      q.tokens[slot] = first4 + b + SeparatorKind(__funnelshift_r(p_lo, p_mid, b) & 0xffu);
      ++slot;
    }
    bits = s1;
    while (bits != 0u) {
      const uint32_t b = __ffs(bits) - 1;
      bits &= bits - 1u;
      B200P_TILE_CHECK(c.error, slot < kTokenQueue)
      q.tokens[slot] = first4 + 32u + b + SeparatorKind(__funnelshift_r(p_mid, p_hi, b) & 0xffu);
      ++slot;
    }
    fill += __shfl_sync(0xffffffffu, incl, 31);
    __syncwarp();

    uint32_t head = 0;
    while (fill - head >= 32u || (done && fill > head)) {
      const uint32_t count = min(fill - head, 32u);
      /* room for one more group in every queue?  (warp-uniform) */
      if (c.n_index + c.n_value > room_pool || c.n_rows > room_rows ||
          (FMT == kLibFM && c.n_index > room_field)) {
        c.overflow = true;
        c.dense = true;
        done = true;
        break;
      }
      const GroupRoles r = TokenRoles<FMT, QID>(tt.global, q.tokens, head, count, lane, carry_pos, carry_first);
      /* every token has at most one role: one store per lane, into the queue of its role (index
         entries grow from the front of the pool, value entries from its back).  The queues are
         addressed as offsets into the warp's one block of 16-bit entries. */
      const uint32_t rel = r.at - tt.win_lo;
      const bool is_index = (r.m_index & me) != 0u, is_value = (r.m_value & me) != 0u;
      const bool is_label = (r.m_label & me) != 0u, is_weight = (r.m_weight & me) != 0u;
      const uint32_t mine = is_index ? r.m_index : (is_value ? r.m_value : (is_label ? r.m_label : r.m_weight));
      const uint32_t rank = __popc(mine & lt);
      const uint32_t at_index = o_pool + c.n_index + rank;
      const uint32_t at_value = o_pool + (prm.cap_pool - 1u - c.n_value) - rank;
      const uint32_t at_label = o_row_at + c.n_rows + rank;
      const uint32_t at_weight = o_weight + c.n_weight + rank;
      const uint32_t slot16 = is_index ? at_index : (is_value ? at_value : (is_label ? at_label : at_weight));
      if (is_index || is_value || is_label || is_weight) q16[slot16] = static_cast<uint16_t>(rel);
      if (is_label) q16[o_row_nnz + c.n_rows + rank] = static_cast<uint16_t>(c.n_index + __popc(r.m_index & lt));
      if (FMT == kLibFM) {
        /* a field id is converted before the reference knows whether the triple is kept
           (libfm_parser.h:73), so a '-' raises even on a token that is then dropped.  Kept field
           tokens are converted later and report it there; only the dropped ones (no index token
           behind them; the group's last token may be one) are looked at here. */
        const uint32_t maybe_dropped = r.m_field & ~(r.m_index >> 1);
        if (maybe_dropped != 0u) {
          if ((maybe_dropped & me) != 0u && tt.byte(r.at) == '-') c.error |= kErrNegativeIndex;
        }
        uint32_t prev_at = __shfl_up_sync(0xffffffffu, r.at, 1);
        if (lane == 0) prev_at = carry_at;
        if (is_index) q16[o_field + c.n_index + rank] = static_cast<uint16_t>(prev_at - tt.win_lo);
      }
      if (FMT == kLibSVM && r.m_qid != 0u) {
        if ((r.m_qid & me) != 0u) {
          q16[o_qid + c.n_qid + __popc(r.m_qid & lt)] = static_cast<uint16_t>(r.qid_at - tt.win_lo);
        }
        c.n_qid += __popc(r.m_qid);
      }
      c.n_rows += __popc(r.m_label);
      c.n_weight += __popc(r.m_weight);
      c.n_index += __popc(r.m_index);
      c.n_value += __popc(r.m_value);
      carry_pos = r.next_pos;
      carry_first = r.next_first;
      carry_at = __shfl_sync(0xffffffffu, r.at, static_cast<int>(count) - 1);
      head += count;
    }
    if (head != 0u && !c.overflow) {
      const uint32_t rest = fill - head;
      uint32_t moved = 0;
      if (lane < rest) moved = q.tokens[head + lane];
      __syncwarp();
      if (lane < rest) q.tokens[lane] = moved;
      __syncwarp();
      fill = rest;
    }
  }
  /* weights and qids live in row-sized queues too; rows <= cap_rows - 32 was checked per group,
     and a row holds at most one of each */
  *out = c;
}

/* ---------------------------------------------------------------- collect: CSV --------------- */
/* A warp owns the fields that start in [lo, hi) (reference csv_parser.h:73-102: every field is a
   float; column label_column is the label, every other column an (index, value) entry). */
__device__ __forceinline__ void CollectFields(const TileParams &prm, const TileText &tt,
                                              const uint8_t *lut, const TileQueues<kCSV> &q,
                                              uint32_t lo, uint32_t hi, int lane, Collected *out) {
  Collected c;
  c.n_rows = c.n_index = c.n_value = c.n_weight = c.n_qid = 0u;
  c.tail = 0u;
  c.has_row = false;
  c.overflow = false;
  c.dense = false;
  c.error = 0u;
  if (lo >= hi) {
    *out = c;
    return;
  }
  const bool labelled = prm.label_column == 0;
  /* class of the byte in front of the sub-range (separator | EOL << 1); the chunk starts a line */
  uint32_t carry = lo == 0u ? 3u : lut[tt.byte(lo - 1u)];
  const uint8_t *const my_win = tt.win + 16u * lane - tt.win_lo;
  uint32_t last_row_nnz = 0;
  for (uint32_t base = lo; base < hi; base += kStepBytes) {
    const uint32_t first_byte = base + 16u * lane;
    const uint4 now = *reinterpret_cast<const uint4 *>(my_win + base);
    const uint32_t cls = ClassifyLut(now, lut, prm.four);
    uint32_t keep = kEven;
    if (base + kStepBytes > hi) keep = Spread16(KeepMask(first_byte, lo, hi));
    uint32_t up = __shfl_up_sync(0xffffffffu, cls >> 30, 1);
    if (lane == 0) up = carry;
    carry = __shfl_sync(0xffffffffu, cls >> 30, 31);
    /* a field starts after a comma or at a line start, unless that byte is an EOL
       (a trailing comma yields no extra column, csv_parser.h:79-95) */
    const uint32_t before = (cls << 2) | up;
    const uint32_t not_eol = ~(cls >> 1) & keep;
    const uint32_t starts = before & not_eol;
    const uint32_t row_starts = (before >> 1) & not_eol;
    const uint32_t values = labelled ? (starts & ~row_starts) : starts;
    const uint32_t nv = __popc(values), nr = __popc(row_starts);
    /* one scan for both counts: values below, rows above bit 16 (a step holds <= 512 of either) */
    const uint32_t both = WarpInclusiveSum(nv | (nr << 16), lane);
    const uint32_t step_values = __shfl_sync(0xffffffffu, both, 31) & 0xffffu;
    const uint32_t step_rows = __shfl_sync(0xffffffffu, both, 31) >> 16;
    if (c.n_rows + step_rows > prm.cap_rows) {   /* cap_pool >= sub_bytes / 2 always holds the fields */
      c.overflow = true;
      c.dense = true;
      break;
    }
    uint32_t vslot = c.n_value + (both & 0xffffu) - nv;
    const uint32_t rel2 = 2u * (first_byte - tt.win_lo);
    uint32_t bits = values;
    while (bits != 0u) {
      const uint32_t b = __ffs(bits) - 1;          /* 2 * byte offset */
      bits &= bits - 1u;
      B200P_TILE_CHECK(c.error, vslot < prm.cap_pool)
      q.pool[vslot] = static_cast<uint16_t>((rel2 + b) >> 1);
      ++vslot;
    }
    if (step_rows != 0u) {
      uint32_t rslot = c.n_rows + (both >> 16) - nr;
      const uint32_t vbase = c.n_value + (both & 0xffffu) - nv;
      bits = row_starts;
      while (bits != 0u) {
        const uint32_t b = __ffs(bits) - 1;
        bits &= bits - 1u;
        q.row_at[rslot] = static_cast<uint16_t>((rel2 + b) >> 1);
        q.row_nnz[rslot] = static_cast<uint16_t>(vbase + __popc(values & ((1u << b) - 1u)));
        ++rslot;
      }
      /* entries in front of the step's last row start */
      const uint32_t who = __ballot_sync(0xffffffffu, nr != 0u);
      const int src = 31 - __clz(who);
      const uint32_t top = 31u - __clz(row_starts | 1u);
      last_row_nnz = __shfl_sync(0xffffffffu, vbase + __popc(values & ((1u << top) - 1u)), src);
      c.has_row = true;
    }
    c.n_value += step_values;
    c.n_rows += step_rows;
  }
  __syncwarp();
  c.n_index = c.n_value;
  c.tail = c.has_row ? c.n_value - last_row_nnz : c.n_value;
  *out = c;
}

/* ---------------------------------------------------------------- the kernel ----------------- */
/* resident CTAs per SM the register budget is held to (shared memory allows as many) */
#ifndef B200P_TILE_CTAS_CSV
#define B200P_TILE_CTAS_CSV 5
#endif
#ifndef B200P_TILE_CTAS_TOKENS
#define B200P_TILE_CTAS_TOKENS 4
#endif
template <int FMT, typename IdxT, bool QID>
__global__ void __launch_bounds__(kTileThreads, FMT == kCSV ? B200P_TILE_CTAS_CSV : B200P_TILE_CTAS_TOKENS)
ParseTiles(const TileParams prm) {
  extern __shared__ __align__(128) uint8_t smem[];
  TileShared &sh = *reinterpret_cast<TileShared *>(smem);
  uint8_t *const window = smem + ((sizeof(TileShared) + 127u) & ~127u);
  const uint32_t window_bytes = TileWindowBytes(prm);
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const uint32_t tile_bytes = kTileWarps * prm.sub_bytes;

  if (threadIdx.x == 0) {
    sh.tile = atomicAdd(reinterpret_cast<unsigned int *>(prm.scratch), 1u);
    sh.flags = 0u;
    BarrierInit(&sh.bar, 1u);
  }
  sh.lut[threadIdx.x] = static_cast<uint8_t>(ClassEntry<FMT>(threadIdx.x));
  __syncthreads();
  const uint32_t tile = sh.tile;
  const uint32_t tile_lo = tile * tile_bytes;
  /* staged bytes: [win_lo, win_hi), of which [win_lo, stage_end) exist in the chunk */
  const uint32_t win_lo = tile == 0u ? 0u : tile_lo - kTileHalo;
  uint8_t *const win = tile == 0u ? window + kTileHalo : window;   /* win[0] <-> win_lo */
  const uint32_t win_hi = tile_lo + tile_bytes + prm.over_bytes;   /* may lie past the chunk: NUL */
  const uint32_t stage_end = min(win_hi, prm.size);
  const uint32_t bulk = (stage_end - win_lo) & ~15u;
  if (threadIdx.x == 0 && bulk != 0u) {
    BarrierExpect(&sh.bar, bulk);
    BulkLoad(win, prm.text + win_lo, bulk, &sh.bar);
  }
  /* the last bytes of a chunk whose size is not a multiple of 16, then NUL up to win_hi */
  for (uint32_t i = win_lo + bulk + threadIdx.x; i < win_hi; i += kTileThreads) {
    win[i - win_lo] = i < prm.size ? prm.text[i] : static_cast<uint8_t>(0);
  }
  if (bulk != 0u) BarrierWait(&sh.bar, 0u);
  __syncthreads();
  (void)window_bytes;

  TileText tt;
  tt.win = win;
  tt.win_lo = win_lo;
  tt.win_hi = win_hi;
  tt.global.data = prm.text;
  tt.global.size = prm.size;

  uint8_t *const queues = window + ((window_bytes + 15u) & ~15u) + warp * TileWarpQueueBytes<FMT>(prm);
  const TileQueues<FMT> q = CarveQueues<FMT>(queues, prm);
  const uint32_t lo = min(tile_lo + warp * prm.sub_bytes, prm.size);
  const uint32_t hi = min(lo + prm.sub_bytes, prm.size);

  /* ---- collect ---- */
  Collected got;
  if constexpr (FMT == kCSV) {
    CollectFields(prm, tt, sh.lut, q, lo, hi, lane, &got);
  } else {
    CollectTokens<FMT, QID>(prm, tt, sh.lut, q, lo, hi, lane, &got);
  }
  if (lane == 0) {
    sh.count[warp][kRows] = got.n_rows;
    sh.count[warp][kNnz] = got.n_index;
    sh.count[warp][kValues] = got.n_value;
    sh.count[warp][kWeights] = got.n_weight;
    sh.count[warp][kQids] = got.n_qid;
    sh.count[warp][5] = got.tail;
    sh.count[warp][6] = got.has_row ? 1u : 0u;
    if (got.overflow) atomicOr(&sh.flags, static_cast<uint32_t>(kErrFallback) | (got.dense ? static_cast<uint32_t>(kErrDense) : 0u));
  }
  __syncthreads();

  /* ---- place: CTA scan of the 8 warps, look-back over the tiles in front ---- */
  if (warp == 0) {
    const bool fallback = (sh.flags & kErrFallback) != 0u;
    TileTotals mine;
#pragma unroll
    for (int c = 0; c < kNumCounters; ++c) mine.c[c] = 0u;
    mine.tail = 0u;
    mine.stop = false;
    if (!fallback) {
      for (int w = 0; w < kTileWarps; ++w) {
#pragma unroll
        for (int c = 0; c < kNumCounters; ++c) mine.c[c] += sh.count[w][c];
        if (sh.count[w][6] != 0u) {
          mine.tail = sh.count[w][5];
          mine.stop = true;
        } else {
          mine.tail += sh.count[w][5];
        }
      }
    }
    unsigned long long *agg = reinterpret_cast<unsigned long long *>(prm.scratch + kTileHeaderBytes);
    unsigned long long *pre = agg + static_cast<size_t>(prm.num_tiles) * kStatusWords;
    if (lane == 0 && tile != 0u) PublishStatus(agg + static_cast<size_t>(tile) * kStatusWords, mine);
    TileTotals excl;
#pragma unroll
    for (int c = 0; c < kNumCounters; ++c) excl.c[c] = 0u;
    excl.tail = 0u;
    excl.stop = true;
    bool ok = true;
    if (tile != 0u) ok = LookBack(prm, tile, lane, &excl);
    if (lane == 0) {
      TileTotals incl;
#pragma unroll
      for (int c = 0; c < kNumCounters; ++c) incl.c[c] = excl.c[c] + mine.c[c];
      incl.tail = mine.stop ? mine.tail : excl.tail + mine.tail;
      incl.stop = true;
      PublishStatus(pre + static_cast<size_t>(tile) * kStatusWords, incl);
      /* where each warp's output starts */
      uint32_t run[kNumCounters];
#pragma unroll
      for (int c = 0; c < kNumCounters; ++c) run[c] = excl.c[c];
      uint32_t open = excl.tail;
      for (int w = 0; w < kTileWarps; ++w) {
#pragma unroll
        for (int c = 0; c < kNumCounters; ++c) {
          sh.base[w][c] = run[c];
          run[c] += sh.count[w][c];
        }
        sh.base[w][5] = open;
        open = sh.count[w][6] != 0u ? sh.count[w][5] : open + sh.count[w][5];
      }
      uint32_t err = fallback ? (sh.flags & (kErrFallback | kErrDense)) : 0u;
      if (!ok) err |= kErrFallback | kErrInternal;
      if (tile + 1u == prm.num_tiles) {
        /* chunk totals, and the RowBlockContainer::GetBlock consistency check (row_block.h:177) */
#pragma unroll
        for (int c = 0; c < kNumCounters; ++c) prm.summary->total[c] = incl.c[c];
        prm.summary->ranges = prm.num_tiles;
        if (incl.c[kValues] != 0u && incl.c[kValues] != incl.c[kNnz]) err |= kErrValueCount;
        if (incl.c[kRows] <= prm.out_rows && incl.c[kNnz] <= prm.out_nnz) {
          prm.offset[incl.c[kRows]] = incl.c[kNnz];
        } else {
          err |= kErrCapacity;
        }
      }
      if (err != 0u) {
        atomicOr(&prm.summary->error, err);
        atomicOr(&sh.flags, err);
      }
    }
  }
  __syncthreads();
  if ((sh.flags & kErrFallback) != 0u) return;   /* this chunk is parsed again by the two-pass path */

  /* ---- convert ---- */
  const uint32_t b_rows = sh.base[warp][kRows], b_nnz = sh.base[warp][kNnz];
  const uint32_t b_value = sh.base[warp][kValues], b_weight = sh.base[warp][kWeights];
  const uint32_t b_qid = sh.base[warp][kQids];
  uint32_t error = got.error;
  if (static_cast<unsigned long long>(b_rows) + got.n_rows > prm.out_rows ||
      static_cast<unsigned long long>(b_nnz) + got.n_index > prm.out_nnz) {
    /* the arrays cannot hold this chunk: the host sees kErrCapacity, grows them, launches again */
    if (lane == 0) atomicOr(&prm.summary->error, static_cast<uint32_t>(kErrCapacity));
    return;
  }
  IdxT *const index = reinterpret_cast<IdxT *>(prm.index);
  TileRelText rt;
  rt.win = win;
  rt.win_bytes = win_hi - win_lo;
  rt.win_lo = win_lo;
  rt.global = tt.global;
  if constexpr (FMT == kCSV) {
    const bool labelled = prm.label_column == 0;
    for (uint32_t i = lane; i < got.n_value; i += 32u) {
      prm.value[b_value + i] = ToFloatFast(rt, q.pool[i]);
    }
    for (uint32_t i = lane; i < got.n_rows; i += 32u) {
      prm.label[b_rows + i] = labelled ? ToFloatFast(rt, q.row_at[i]) : 0.0f;
      prm.offset[b_rows + i] = static_cast<unsigned long long>(b_nnz + q.row_nnz[i]);
    }
    /* column ids: the j-th value of a row has id j (csv_parser.h:84-88).  Values in front of the
       warp's first row start continue the row that was open when the sub-range began. */
    const uint32_t open = sh.base[warp][5];
    if (got.n_value >= 16u * got.n_rows) {
      const uint32_t head_len = got.n_rows != 0u ? q.row_nnz[0] : got.n_value;
      for (uint32_t j = lane; j < head_len; j += 32u) index[b_nnz + j] = static_cast<IdxT>(open + j);
      for (uint32_t r = 0; r < got.n_rows; ++r) {
        const uint32_t s = q.row_nnz[r];
        const uint32_t e = r + 1u < got.n_rows ? q.row_nnz[r + 1u] : got.n_value;
        for (uint32_t j = s + lane; j < e; j += 32u) index[b_nnz + j] = static_cast<IdxT>(j - s);
      }
    } else {
      /* short rows: every value finds its row by bisection over the row starts */
      for (uint32_t i = lane; i < got.n_value; i += 32u) {
        uint32_t a = 0, b = got.n_rows;     /* rows [0, a) start at or before value i */
        while (a < b) {
          const uint32_t m = (a + b) >> 1;
          if (q.row_nnz[m] <= i) a = m + 1u; else b = m;
        }
        index[b_nnz + i] = static_cast<IdxT>(a == 0u ? open + i : i - q.row_nnz[a - 1u]);
      }
    }
  } else {
    for (uint32_t i = lane; i < got.n_rows; i += 32u) {
      prm.label[b_rows + i] = ToFloatFast(rt, q.row_at[i]);
      prm.offset[b_rows + i] = static_cast<unsigned long long>(b_nnz + q.row_nnz[i]);
    }
    for (uint32_t i = lane; i < got.n_weight; i += 32u) {
      prm.weight[b_weight + i] = ToFloatFast(rt, q.weight[i]);
    }
    if constexpr (FMT == kLibSVM) {
      for (uint32_t i = lane; i < got.n_qid; i += 32u) {
        prm.qid[b_qid + i] = ToQid(tt.global, win_lo + q.qid[i]);
      }
    }
    bool neg = false;
    for (uint32_t i = lane; i < got.n_index; i += 32u) {
      index[b_nnz + i] = static_cast<IdxT>(ToUintFast<sizeof(IdxT) == 8>(rt, q.pool[i], &neg));
      if constexpr (FMT == kLibFM) {
        reinterpret_cast<IdxT *>(prm.field)[b_nnz + i] =
            static_cast<IdxT>(ToUintFast<sizeof(IdxT) == 8>(rt, q.field[i], &neg));
      }
    }
    if (neg) error |= kErrNegativeIndex;
    for (uint32_t i = lane; i < got.n_value; i += 32u) {
      prm.value[b_value + i] = ToFloatFast(rt, q.pool[prm.cap_pool - 1u - i]);
    }
  }
  error = __reduce_or_sync(0xffffffffu, error);
  if (error != 0u && lane == 0) atomicOr(&prm.summary->error, error);
}

}  // namespace b200p
#endif  // DMLC_CORE_B200_TILE_KERNELS_CUH_
