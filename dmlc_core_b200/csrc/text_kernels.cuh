/*
 * sm_100a kernels that turn one chunk of LibSVM / LibFM / CSV text, resident in HBM, into the
 * CSR arrays of a dmlc RowBlock.  They replace the CPU work of
 *   TextParserBase::FillData         reference src/data/text_parser.h:93-134
 *   LibSVMParser::ParseBlock         reference src/data/libsvm_parser.h:35-99
 *   LibFMParser::ParseBlock          reference src/data/libfm_parser.h:35-93
 *   CSVParser::ParseBlock            reference src/data/csv_parser.h:63-104
 * and are bit-exact with them (oracle: oracle/dmlc_oracle.c, oracle/_ref).
 *
 * Shape of the computation (byte and integer work, no tensor cores; DRAM traffic equals the
 * algorithmic bytes, the measured limit is instruction issue -- DESIGN.md section 6):
 *
 *   pass 1  WalkText<..., EMIT=false>   count rows / nnz / values / weights / qids per byte range
 *   scan    ScanTiles + ScanFinish      exclusive prefix over ranges -> where each range writes
 *   pass 2  WalkText<..., EMIT=true>    same walk, now converting tokens and storing CSR arrays
 *
 * The chunk is cut into fixed byte ranges; a warp owns the lines that START inside its range
 * (the reference slices a chunk the same way for its OpenMP team, text_parser.h:104-118), so no
 * parsing state ever crosses a range.  Inside a range the warp streams 512-byte steps: every
 * lane loads 16 contiguous bytes (one coalesced 512 B request per warp) and classifies them
 * through a 256-byte table in shared memory (PRMT + LDS.U8 + IMAD per byte: the logic/shift ALU
 * pipe is what bounds these kernels, so the work is spread over the LSU and FMA pipes) into
 * interleaved class masks; the starts of digit runs ("tokens"; CSV: field starts) are compacted
 * into a per-warp queue in shared memory, each tagged with the kind of separator in front of
 * it.  Tokens are then handled 32 at a time, one per lane: warp ballots + popc/clz turn "bound
 * by ':' to the previous token" / "starts a line" into chain positions, roles (label / weight /
 * qid / field / index / value), row ids and output slots -- the exact pairing rules of
 * ParsePair / ParseTriple (reference src/data/strtonum.h:227-303); only unusual separators are
 * walked byte by byte.  Tokens of one role are queued again so that conversion always runs with
 * 32 busy lanes doing the same arithmetic, and every store is 32 consecutive elements of one
 * CSR array.
 */
#ifndef DMLC_CORE_B200_TEXT_KERNELS_CUH_
#define DMLC_CORE_B200_TEXT_KERNELS_CUH_

#include <cuda_runtime.h>
#include <stdint.h>

#include "strtonum_device.h"

namespace b200p {

enum Format { kLibSVM = 0, kLibFM = 1, kCSV = 2 };

/* counters carried per range through the scan; order fixed (index into arrays below) */
enum Counter { kRows = 0, kNnz = 1, kValues = 2, kWeights = 3, kQids = 4, kNumCounters = 5 };

/* error bits surfaced to the host (-> dmlc::Error in the C++ layer) */
enum ErrorBits {
  kErrNegativeIndex = 1u,   /* reference strtonum.h:141  CHECK_EQ(sign, true) */
  kErrValueCount = 2u,      /* reference row_block.h:177 offset.back()==value.size()||empty */
  kErrCapacity = 4u,        /* output buffers too small: host grows them and re-runs pass 2 */
  kErrInternal = 8u,        /* checked build only: a queue or output index left its bounds */
  kErrFallback = 16u,       /* one-launch kernel (tile_kernels.cuh) gave up: parse the chunk with the two-pass walk */
  kErrDense = 32u,          /* ... because a sub-range holds more tokens or rows than its queues (not: a long line) */
};

/*
 * Checked build (-DB200P_CHECKED, libdmlc_b200_checked.so): every shared-memory queue write and
 * every store into a CSR array is bounds-checked against the queue size / the totals pass 1
 * counted; a violation sets kErrInternal instead of writing.  compute-sanitizer is not available
 * on the GPU pool, so the GPU tests run a corpus + fuzz subset through this build instead
 * (tests/test_gpu_checked_build.py).  The production build compiles the checks away.
 */
#ifdef B200P_CHECKED
#define B200P_CHECK(prm, cond) \
  if (!(cond)) { atomicOr(&(prm).summary->error, static_cast<unsigned int>(kErrInternal)); } else
#else
#define B200P_CHECK(prm, cond)
#endif

constexpr int kStepBytes = 512;        /* one warp step: 32 lanes x 16 B */
#ifndef B200P_WARPS_PER_CTA   /* tuning experiments only */
#define B200P_WARPS_PER_CTA 8
#endif
constexpr int kWarpsPerCta = B200P_WARPS_PER_CTA;
constexpr int kCtaThreads = kWarpsPerCta * 32;
constexpr int kTokenQueue = 320;       /* < 32 left over + <= 256 new digit runs per step */
constexpr int kFieldQueue = 576;       /* CSV: < 32 left over + <= 512 field starts per step */
constexpr int kScanTile = 1024;         /* ranges scanned by one CTA of the prefix-sum kernels */
constexpr int kRoleQueue = 64;         /* < 32 left over + <= 32 new per token group */

/* Result header written by ScanFinish (device memory, copied to the host per chunk). */
struct ChunkSummary {
  unsigned long long total[kNumCounters];
  unsigned int error;
  unsigned int ranges;
};

struct WalkParams {
  const uint8_t *text;     /* 16-byte aligned */
  uint32_t size;           /* bytes, < 2^31 */
  uint32_t range_bytes;    /* multiple of kStepBytes */
  uint32_t num_ranges;
  int32_t label_column;    /* CSV */
  uint32_t sixteen;        /* always 16, for the same reason as `four` below */
  uint32_t four;           /* always 4. Passed in so that the compiler cannot fold `acc * 4 + e`
                              into a shift-add on the (saturated) ALU pipe: as a run-time value
                              it stays an IMAD on the otherwise idle FMA pipe. */
  uint32_t *counts;        /* [kNumCounters][num_ranges]: pass 1 writes counts, the scan rewrites
                              them in place as exclusive prefixes within tiles of kScanTile */
  uint32_t *tile_base;     /* [kNumCounters][num_tiles]: exclusive prefix of the tile totals */
  uint32_t num_tiles;
  ChunkSummary *summary;
  /* outputs (pass 2) and their capacities in elements */
  unsigned long long *offset;
  float *label;
  float *weight;
  unsigned long long *qid;
  void *field;
  void *index;
  float *value;
  unsigned long long cap_rows;
  unsigned long long cap_nnz;
};

/* ------------------------------------------------------------- SWAR end-of-line scan -------- */
/* Used only by FindEol (the short forward look for a range's first line start); the walks
   themselves classify bytes through the shared-memory table below.
   xm = word & 0x7f7f7f7f; bit 7 of each result byte is the predicate; bytes >= 0x80 are removed
   with & ~word. */
__device__ __forceinline__ uint32_t ByteEq(uint32_t xm, uint32_t c) {
  return ~((xm ^ (c * 0x01010101u)) + 0x7f7f7f7fu);
}
/* bit 7 of each byte -> 4 adjacent bits */
__device__ __forceinline__ uint32_t Gather4(uint32_t m) {
  return (((m & 0x80808080u) >> 7) * 0x00204081u >> 21) & 0xFu;
}
__device__ __forceinline__ uint32_t WordEol(uint32_t w) {
  const uint32_t xm = w & 0x7f7f7f7fu;
  return (ByteEq(xm, '\n') | ByteEq(xm, '\r')) & ~w;
}

/*
 * Byte classes through a 256-byte table in shared memory.  The SWAR classifier costs ~50 logic
 * instructions per 16 bytes, all on the half-rate ALU pipe that bounds these kernels; the table
 * walk costs one PRMT (ALU), one LDS.U8 (LSU) and one IMAD (FMA pipe) per byte.  ASCII bytes map
 * to distinct banks or to the same word (broadcast), so the lookups are conflict-free.
 *
 * Table entry: bit 0 = class A (LibSVM/LibFM: digit-run byte; CSV: ',' or EOL), bit 1 = EOL.
 * Result: a 32-bit interleaved mask, bit 2j = A of byte j, bit 2j+1 = EOL of byte j.
 */
constexpr uint32_t kEven = 0x55555555u;

template <int FMT>
__device__ __forceinline__ uint32_t ClassEntry(uint32_t byte) {
  const bool eol = byte == '\n' || byte == '\r';
  const bool a = FMT == kCSV ? (eol || byte == ',') : IsDigitChar(byte);
  uint32_t e = (a ? 1u : 0u) | (eol ? 2u : 0u);
  if (FMT != kCSV) e |= (byte == ':' ? 4u : 0u) | (IsBlank(byte) ? 8u : 0u);
  return e;
}

/* LibSVM / LibFM: four classes per byte (digit-run 1, EOL 2, ':' 4, blank 8), one nibble per byte:
   m0 = bytes 0..7, m1 = bytes 8..15 of the lane's 16. */
constexpr uint32_t kNibD = 0x11111111u;
__device__ __forceinline__ void ClassifyLut4(const uint4 v, const uint8_t *lut, uint32_t sixteen,
                                             uint32_t *m0, uint32_t *m1) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
  uint32_t lo = 0, hi = 0;
#pragma unroll
  for (int j = 7; j >= 0; --j) {
    lo = lo * sixteen + lut[__byte_perm(w[j >> 2], 0u, 0x4440u | (j & 3))];
    hi = hi * sixteen + lut[__byte_perm(w[2 + (j >> 2)], 0u, 0x4440u | (j & 3))];
  }
  *m0 = lo;
  *m1 = hi;
}

/* 8-bit byte mask -> 0xF in the nibble of every selected byte */
__device__ __forceinline__ uint32_t SpreadNibbles8(uint32_t m) {
  m = (m | (m << 12)) & 0x000f000fu;
  m = (m | (m << 6)) & 0x03030303u;
  m = (m | (m << 3)) & 0x11111111u;
  return m * 15u;
}

/*
 * Separator kind of a token from the classes of the two bytes in front of it (c = class nibble of
 * byte -2 in the low half, of byte -1 in the high half):
 *   3 line start   the byte before is an EOL
 *   1 bound        "<digit-run byte>:"   -- separator is exactly ':'
 *   2 unbound      "<digit-run byte><blank>" -- separator is exactly one blank
 *   0 general      anything else: TokenGroup walks back over the separator
 * Index = (EOL, ':', blank of byte -1) | digit-run of byte -2 << 3; two bits per entry.
 */
__device__ __forceinline__ uint32_t SeparatorKind(uint32_t c) {
  const uint32_t idx = ((c >> 5) & 7u) | ((c & 1u) << 3);
  /* idx: bit0 EOL, bit1 colon, bit2 blank, bit3 previous-previous is a digit-run byte.
     EOL -> 3 regardless; colon&D -> 1; blank&D -> 2 */
  constexpr uint32_t kTable =
      (3u << 2) | (3u << 6) | (3u << 10) | (3u << 14) |            /* idx 1,3,5,7: EOL */
      (3u << 18) | (1u << 20) | (3u << 22) | (2u << 24) | (3u << 26) | (3u << 30); /* idx 9..15 */
  return (kTable >> (2u * idx)) & 3u;
}

__device__ __forceinline__ uint32_t ClassifyLut(const uint4 v, const uint8_t *lut, uint32_t four) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
  uint32_t lo = 0, hi = 0; /* two Horner chains: bytes 7..0 and 15..8 */
#pragma unroll
  for (int j = 7; j >= 0; --j) {
    lo = lo * four + lut[__byte_perm(w[j >> 2], 0u, 0x4440u | (j & 3))];
    hi = hi * four + lut[__byte_perm(w[2 + (j >> 2)], 0u, 0x4440u | (j & 3))];
  }
  return lo | (hi << 16);
}

/* 16-bit byte mask -> the even bits of a 32-bit interleaved mask */
__device__ __forceinline__ uint32_t Spread16(uint32_t m) {
  m = (m | (m << 8)) & 0x00ff00ffu;
  m = (m | (m << 4)) & 0x0f0f0f0fu;
  m = (m | (m << 2)) & 0x33333333u;
  m = (m | (m << 1)) & 0x55555555u;
  return m;
}

/* 16 bytes at `pos` (multiple of 16); bytes at or past `size` read as NUL */
__device__ __forceinline__ uint4 Load16(const uint8_t *text, uint32_t size, uint32_t pos) {
  if (pos + 16u <= size) {
    return __ldg(reinterpret_cast<const uint4 *>(text + pos));
  }
  uint32_t w[4] = {0u, 0u, 0u, 0u};
  for (uint32_t i = 0; i < 16u; ++i) {
    if (pos + i < size) w[i >> 2] |= static_cast<uint32_t>(text[pos + i]) << ((i & 3u) * 8u);
  }
  return make_uint4(w[0], w[1], w[2], w[3]);
}

/* bits [lo, hi) of a lane's 16-bit mask, for absolute byte bounds against the lane's first byte */
__device__ __forceinline__ uint32_t KeepMask(uint32_t lane_first, uint32_t from, uint32_t to) {
  const uint32_t lo = from > lane_first ? min(from - lane_first, 16u) : 0u;
  const uint32_t hi = to > lane_first ? min(to - lane_first, 16u) : 0u;
  return ((1u << hi) - 1u) & ~((1u << lo) - 1u);
}

__device__ __forceinline__ uint32_t LaneMaskLt() {
  uint32_t m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}
__device__ __forceinline__ uint32_t LaneMaskLe() {
  uint32_t m;
  asm("mov.u32 %0, %%lanemask_le;" : "=r"(m));
  return m;
}

/* inclusive warp prefix sum of small non-negative ints */
__device__ __forceinline__ uint32_t WarpInclusiveSum(uint32_t v, int lane) {
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t o = __shfl_up_sync(0xffffffffu, v, d);
    if (lane >= d) v += o;
  }
  return v;
}

/* first end-of-line byte in [from, to): returns its position, or 0xffffffff */
__device__ uint32_t FindEol(const uint8_t *text, uint32_t size, uint32_t from, uint32_t to,
                            int lane) {
  if (to > size) to = size;
  for (uint32_t base = from & ~uint32_t(kStepBytes - 1); base < to; base += kStepBytes) {
    const uint32_t first = base + 16u * lane;
    const uint4 v = Load16(text, size, first);
    uint32_t e = Gather4(WordEol(v.x)) | (Gather4(WordEol(v.y)) << 4) |
                 (Gather4(WordEol(v.z)) << 8) | (Gather4(WordEol(v.w)) << 12);
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

/* ------------------------------------------------------------------ per-warp scratch -------- */
template <int FMT>
struct WarpScratch {
  uint32_t tokens[FMT == kCSV ? kFieldQueue : kTokenQueue];
  /* role queues (pass 2): LibSVM/LibFM use label, weight, qid, index, value (+ field);
     CSV converts straight out of the token queue and uses none of them */
  uint32_t q_label[FMT == kCSV ? 1 : kRoleQueue];
  uint32_t q_weight[FMT == kCSV ? 1 : kRoleQueue];
  uint32_t q_qid[FMT == kLibSVM ? kRoleQueue : 1];
  uint32_t q_field[FMT == kLibFM ? kRoleQueue : 1];
  uint32_t q_index[FMT == kCSV ? 1 : kRoleQueue];
  uint32_t q_value[FMT == kCSV ? 1 : kRoleQueue];
};

/* State of one warp while it walks its range. All members are warp-uniform except error. */
template <typename IdxT>
struct WalkState {
  uint32_t seen[kNumCounters];  /* pass 1: counts so far; pass 2: absolute output cursors for
                                   the *role pass* (offset/label row ids, nnz ids) */
  /* chain context of the last token handled (ParsePair/ParseTriple pairing) */
  int carry_pos;        /* position of that token in its ':'-bound chain (1-based) */
  bool carry_first;     /* chain is the first one on its line (label[:weight]...) */
  uint32_t carry_at;    /* byte position of that token */
  uint32_t carry_col;   /* CSV: column of the last field handled */
  /* pass 2: role queues fill levels and how many elements of each array were stored */
  uint32_t n_label, n_weight, n_qid, n_index, n_value;
  uint32_t w_label, w_weight, w_qid, w_index, w_value;
  uint32_t error;
};

template <typename T>
__device__ __forceinline__ T MaxOf(T a, T b) {
  return a > b ? a : b;
}
template <typename IdxT>
__device__ __forceinline__ void StoreIdx(void *base, uint32_t at, uint64_t v) {
  reinterpret_cast<IdxT *>(base)[at] = static_cast<IdxT>(v);
}

/* Drain helper: lanes [0, take) pop queue entries and run `conv`; the rest of the queue moves
   to the front. `take` is 32 in steady state, smaller only at the final flush. */
__device__ __forceinline__ void QueuePop(uint32_t *q, uint32_t &fill, uint32_t take, int lane) {
  const uint32_t rest = fill - take;
  uint32_t moved = 0;
  if (lane < rest) moved = q[take + lane];
  __syncwarp();
  if (lane < rest) q[lane] = moved;
  __syncwarp();
  fill = rest;
}

template <int FMT, typename IdxT>
__device__ __forceinline__ void DrainRoleQueues(const WalkParams &prm, const Text &txt,
                                                WarpScratch<FMT> &ws, WalkState<IdxT> &st, int lane,
                                                bool flush) {
  const uint32_t need = flush ? 1u : 32u;
  while (st.n_label >= need) {
    const uint32_t take = min(st.n_label, 32u);
    if (lane < take) {
      B200P_CHECK(prm, st.w_label + lane < prm.summary->total[kRows])
      prm.label[st.w_label + lane] = ToFloatFast(txt, ws.q_label[lane]);
    }
    st.w_label += take;
    QueuePop(ws.q_label, st.n_label, take, lane);
  }
  while (st.n_weight >= need) {
    const uint32_t take = min(st.n_weight, 32u);
    if (lane < take) {
      B200P_CHECK(prm, st.w_weight + lane < prm.summary->total[kWeights])
      prm.weight[st.w_weight + lane] = ToFloatFast(txt, ws.q_weight[lane]);
    }
    st.w_weight += take;
    QueuePop(ws.q_weight, st.n_weight, take, lane);
  }
  if (FMT == kLibSVM) {
    while (st.n_qid >= need) {
      const uint32_t take = min(st.n_qid, 32u);
      if (lane < take) {
        B200P_CHECK(prm, st.w_qid + lane < prm.summary->total[kQids])
        prm.qid[st.w_qid + lane] = ToQid(txt, ws.q_qid[lane]);
      }
      st.w_qid += take;
      QueuePop(ws.q_qid, st.n_qid, take, lane);
    }
  }
  while (st.n_index >= need) {
    const uint32_t take = min(st.n_index, 32u);
    bool neg = false;
    if (lane < take) {
      B200P_CHECK(prm, st.w_index + lane < prm.summary->total[kNnz]) {
        StoreIdx<IdxT>(prm.index, st.w_index + lane,
                       ToUintFast<sizeof(IdxT) == 8>(txt, ws.q_index[lane], &neg));
        if (FMT == kLibFM) {
          StoreIdx<IdxT>(prm.field, st.w_index + lane,
                         ToUintFast<sizeof(IdxT) == 8>(txt, ws.q_field[lane], &neg));
        }
      }
    }
    if (neg) st.error |= kErrNegativeIndex; /* lane-local; OR-reduced at the end of the walk */
    st.w_index += take;
    if (FMT == kLibFM) {
      uint32_t fill = st.n_index;
      QueuePop(ws.q_field, fill, take, lane);
    }
    QueuePop(ws.q_index, st.n_index, take, lane);
  }
  while (st.n_value >= need) {
    const uint32_t take = min(st.n_value, 32u);
    if (lane < take) {
      B200P_CHECK(prm, st.w_value + lane < prm.summary->total[kValues])
      prm.value[st.w_value + lane] = ToFloatFast(txt, ws.q_value[lane]);
    }
    st.w_value += take;
    QueuePop(ws.q_value, st.n_value, take, lane);
  }
}

/*
 * One group of <= 32 LibSVM / LibFM tokens (digit runs), one per lane, in file order.
 * Pairing rules restated from ParsePair / ParseTriple (reference strtonum.h:227-303):
 * a token is "bound" to the previous one when the separator between them is
 * [blank]* ':' [anything but digitchars / EOL]*; bound tokens form a chain.  In the first chain
 * of a line positions 1,2 are label, weight (libsvm_parser.h:50-61); everything after that is
 * consumed in groups of 2 (index[:value]) or 3 (field:index[:value]).
 */
/* true when the six bytes in front of position `at` are "<digit-run byte> qid:" -- the one longer
   separator well-formed ranking data has on every row (libsvm_parser.h:66-75).  Two aligned word
   pairs instead of six byte loads behind short-circuit tests: one trip to memory, not six. */
__device__ __forceinline__ bool QidBehind(const Text &txt, uint32_t at) {
  if (at < 8u || ((at - 8u) & ~3u) + 12u > txt.size) {
    if (at < 6u) return false;
    const uint32_t c1 = txt.at(at - 1u), c2 = txt.at(at - 2u), c3 = txt.at(at - 3u), c4 = txt.at(at - 4u),
                   c5 = txt.at(at - 5u), c6 = txt.at(at - 6u);
    return c1 == ':' && c2 == 'd' && c3 == 'i' && c4 == 'q' && c5 == ' ' && IsDigitChar(c6);
  }
  const uint32_t p = at - 8u, a = p & ~3u, k = p & 3u;
  const uint32_t w0 = Load4(txt.data + a), w1 = Load4(txt.data + a + 4), w2 = Load4(txt.data + a + 8);
  const uint32_t x0 = BytesFrom(w0, w1, k), x1 = BytesFrom(w1, w2, k);   /* bytes at-8..at-5, at-4..at-1 */
  return x1 == 0x3a646971u /* "qid:" */ && (x0 >> 24) == ' ' && IsDigitChar((x0 >> 16) & 0xffu);
}

struct GroupRoles {
  uint32_t m_label, m_weight, m_index, m_value, m_qid, m_field; /* bit i = token i of the group */
  uint32_t at;          /* this lane's token: byte position */
  uint32_t qid_at;      /* this lane's token, when it is a qid number: the byte after "qid:" */
  int next_pos;         /* chain context handed to the next group */
  bool next_first;
};

template <int FMT, bool QID>
__device__ __forceinline__ GroupRoles TokenRoles(const Text &txt, const uint32_t *tokens,
                                                 uint32_t head, uint32_t count, int lane,
                                                 int carry_pos, bool carry_first) {
  const bool active = lane < count;
  const uint32_t entry = active ? tokens[head + lane] : 3u;
  const uint32_t at = entry >> 2;
  const uint32_t kind = entry & 3u;   /* separator kind, see SeparatorKind() */

  const bool general = active && kind == 0u;
  const uint32_t general_m = __ballot_sync(0xffffffffu, general);
  /* QID instantiation (chosen by the host once a chunk has reported qids): the one longer
     separator that ranking data has on every row, "<digit-run byte> qid:" (libsvm_parser.h:66-75),
     is recognised by its six bytes, and a group whose only unusual separators are of that form
     keeps the mask arithmetic below.  The plain instantiation sends such groups down the general
     path; both give the same result (the extra registers cost the qid-free shapes 3-4 %, hence
     two instantiations). */
  bool qid_number = false;
  uint32_t qid_numbers = 0u;
  if (QID && FMT == kLibSVM && general_m != 0u) {
    if (general) qid_number = QidBehind(txt, at);
    qid_numbers = __ballot_sync(0xffffffffu, qid_number);
  }

  /* role masks of the group (bit i = token i) and the chain context handed to the next group */
  uint32_t m_label, m_weight, m_index, m_value, m_qid = 0u, m_field = 0u;
  int next_pos;
  bool next_first;
  int prev_pos = 0;   /* general path only: chain position of the token before this one */
  bool simple = false;
  uint32_t qid_at = 0;   /* byte after "qid:" when this token follows the probe */

  if (general_m == qid_numbers) {
    /* Common case.  With only the three simple separator kinds in the group and no chain longer
       than three tokens, chain positions and roles are plain mask arithmetic on warp-uniform
       ballots -- no per-lane chain walk.  Longer chains and unusual separators take the general
       code below; both implement the same rules (the checked build verifies that pass 2
       produces exactly what pass 1 counted, range by range). */
    const uint32_t act_m = 0xffffffffu >> (32u - count);   /* count is 1..32 */
    const uint32_t line_m = __ballot_sync(0xffffffffu, active && kind == 3u);
    const uint32_t bound_m = __ballot_sync(0xffffffffu, active && kind == 1u);
    const uint32_t start_m = act_m & ~bound_m;                    /* chain position 1 */
    const uint32_t c1 = carry_pos == 1 ? 1u : 0u, c2 = carry_pos == 2 ? 1u : 0u;
    const uint32_t c3 = carry_pos >= 3 ? 1u : 0u, cf = carry_first ? 1u : 0u;
    const uint32_t p2 = bound_m & ((start_m << 1) | c1);          /* chain position 2 */
    const uint32_t p3 = bound_m & ((p2 << 1) | c2);               /* chain position 3 */
    const uint32_t p4 = bound_m & ((p3 << 1) | c3);               /* longer chains */
    if (p4 == 0u) {
      simple = true;
      const uint32_t p2f = p2 & ((line_m << 1) | (c1 & cf));      /* ... in the line's first chain */
      const uint32_t p3f = p3 & ((p2f << 1) | (c2 & cf));
      m_label = line_m;
      m_weight = p2f;
      if (FMT == kLibSVM) {
        m_index = (start_m & ~line_m) | p3;   /* first chain: label, weight, index; else index, value, index */
        m_value = p2 & ~p2f;
        if (QID) {
          /* a qid number counts when it follows the label or label:weight; it is read as an
             index as well (the reference rewinds to it, libsvm_parser.h:74) */
          m_qid = qid_numbers & (((line_m | p2f) << 1) | (cf & (c1 | c2)));
          qid_at = at;
        }
      } else {
        m_field = (start_m & ~line_m) | p3f;  /* field candidates (converted even when dropped) */
        m_index = p2 & ~p2f;                  /* field:index[:value] */
        m_value = p3 & ~p3f;
      }
      const uint32_t last = 1u << (count - 1u);
      next_pos = (p3 & last) ? 3 : ((p2 & last) ? 2 : 1);
      next_first = ((line_m | p2f | p3f) & last) != 0u;
    }
  }

  if (!simple) {
    /* ---- the separator in front of this token ---- */
    bool line_start = kind == 3u;  /* separator contains an EOL (or the chunk starts here) */
    bool bound = kind == 1u;
    bool qid_a = false, qid_b = false;     /* separator is [blank]*"qid:"... / [space]*"qid:"... */
    uint32_t qid_at_a = 0, qid_at_b = 0;   /* byte after that "qid:" */
    if (general_m != 0u) {
      /* the one longer separator that well-formed ranking data has on every row,
         "<digit-run byte> qid:" (libsvm_parser.h:66-75), is recognised without the walk */
      bool walk = general;
      if (FMT == kLibSVM && general) {
        if (QID ? qid_number : QidBehind(txt, at)) {
          line_start = false;
          bound = false;            /* first non-blank of the separator is 'q' */
          qid_a = qid_b = true;     /* one space, then "qid:": both probe variants match */
          qid_at_a = qid_at_b = at;
          walk = false;
        }
      }
      if (walk) {
        /* anything else: walk back over the separator */
        line_start = true;
        const uint32_t kQid = (uint32_t(':') << 24) | (uint32_t('d') << 16) | (uint32_t('i') << 8) |
                              uint32_t('q');
        uint32_t first_nonblank = 0, window = 0;
        bool in_a = false, in_b = false;
        uint32_t q = at;
        for (;;) {
          if (q == 0u) break;               /* start of chunk: a line starts here */
          --q;
          const uint32_t c = txt.at(q);
          if (IsEol(c)) break;
          if (IsDigitChar(c)) {             /* previous token ends at q */
            line_start = false;
            bound = (first_nonblank == ':');
            if (FMT == kLibSVM) {
              if (!in_a) { qid_a = (window == kQid); qid_at_a = q + 5u; }
              if (!in_b) { qid_b = (window == kQid); qid_at_b = q + 5u; }
            }
            break;
          }
          const bool space = (c == ' ');
          const bool blank = space || c == '\t';
          if (!blank) first_nonblank = c;
          if (FMT == kLibSVM) {
            /* leading-separator tracking for "qid:" (libsvm_parser.h:66-75): after `label` the
               reference has skipped blanks, after `label:weight` only spaces */
            if (blank) {
              if (!in_a) { qid_a = (window == kQid); qid_at_a = q + 5u; in_a = true; }
            } else {
              in_a = false; qid_a = false;
            }
            if (space) {
              if (!in_b) { qid_b = (window == kQid); qid_at_b = q + 5u; in_b = true; }
            } else {
              in_b = false; qid_b = false;
            }
            window = (window << 8) | c;
          }
        }
      }
    }

    /* ---- chain position and role, by ballot ---- */
    const uint32_t act_m = __ballot_sync(0xffffffffu, active);
    const uint32_t nl_m = __ballot_sync(0xffffffffu, active && line_start);
    const uint32_t bd_m = __ballot_sync(0xffffffffu, active && bound);
    const uint32_t starts = act_m & ~bd_m;              /* tokens that open a chain */
    const uint32_t below = starts & LaneMaskLe();
    int pos;
    bool first;
    if (below != 0u) {
      const int j = 31 - __clz(below);
      pos = lane - j + 1;
      first = (nl_m >> j) & 1u;
    } else {
      pos = carry_pos + lane + 1;
      first = carry_first;
    }
    prev_pos = __shfl_up_sync(0xffffffffu, pos, 1);
    int prev_first = __shfl_up_sync(0xffffffffu, static_cast<int>(first), 1);
    if (lane == 0) {
      prev_pos = carry_pos;
      prev_first = carry_first;
    }
    bool is_label = false, is_weight = false, is_index = false, is_value = false, is_field = false;
    if (active) {
      if (first && pos == 1) {
        is_label = true;
      } else if (first && pos == 2) {
        is_weight = true;
      } else {
        const int k = first ? pos - 3 : pos - 1;
        if (FMT == kLibSVM) {
          is_index = (k & 1) == 0;
          is_value = !is_index;
        } else {
          const int r = k % 3;
          is_field = r == 0;   /* emitted only if an index follows: done by the index token */
          is_index = r == 1;
          is_value = r == 2;
        }
      }
    }
    bool is_qid = false;
    if (FMT == kLibSVM) {
      is_qid = active && !line_start && !bound && prev_first && prev_pos <= 2 &&
               (prev_pos == 1 ? qid_a : qid_b);
      qid_at = (prev_pos == 1) ? qid_at_a : qid_at_b;
    }
    m_label = __ballot_sync(0xffffffffu, is_label);
    m_weight = __ballot_sync(0xffffffffu, is_weight);
    m_index = __ballot_sync(0xffffffffu, is_index);
    m_value = __ballot_sync(0xffffffffu, is_value);
    if (FMT == kLibSVM) m_qid = __ballot_sync(0xffffffffu, is_qid);
    if (FMT == kLibFM) m_field = __ballot_sync(0xffffffffu, is_field);
    const int last = static_cast<int>(count) - 1;
    next_pos = __shfl_sync(0xffffffffu, pos, last);
    next_first = __shfl_sync(0xffffffffu, static_cast<int>(first), last) != 0;
  }

  GroupRoles r;
  r.m_label = m_label;
  r.m_weight = m_weight;
  r.m_index = m_index;
  r.m_value = m_value;
  r.m_qid = m_qid;
  r.m_field = m_field;
  r.at = at;
  r.qid_at = qid_at;
  r.next_pos = next_pos;
  r.next_first = next_first;
  return r;
}

/* The roles of one group, then (EMIT) into the role queues of the two-pass walk. */
template <int FMT, typename IdxT, bool EMIT, bool QID>
__device__ __forceinline__ void TokenGroup(const WalkParams &prm, const Text &txt,
                                           WarpScratch<FMT> &ws, WalkState<IdxT> &st, uint32_t head,
                                           uint32_t count, int lane) {
  const GroupRoles roles = TokenRoles<FMT, QID>(txt, ws.tokens, head, count, lane, st.carry_pos,
                                                st.carry_first);
  const uint32_t m_label = roles.m_label, m_weight = roles.m_weight, m_index = roles.m_index;
  const uint32_t m_value = roles.m_value, m_qid = roles.m_qid, m_field = roles.m_field;
  const uint32_t at = roles.at, qid_at = roles.qid_at;
  const int next_pos = roles.next_pos;
  const bool next_first = roles.next_first;

  if (EMIT) {
    const uint32_t lt = LaneMaskLt();
    const uint32_t me = 1u << lane;
    const bool is_label = (m_label & me) != 0u, is_weight = (m_weight & me) != 0u;
    const bool is_index = (m_index & me) != 0u, is_value = (m_value & me) != 0u;
    /* row pointer: a label token opens row r; offset[r] = features stored before it
       (libsvm_parser.h:62-64 push the running index count) */
    if (is_label) {
      B200P_CHECK(prm, st.seen[kRows] + __popc(m_label & lt) < prm.summary->total[kRows])
      prm.offset[st.seen[kRows] + __popc(m_label & lt)] =
          static_cast<unsigned long long>(st.seen[kNnz] + __popc(m_index & lt));
    }
    if (FMT == kLibFM && (m_field & me) != 0u && txt.at(at) == '-') {
      /* a field id is converted before the reference knows whether the triple is kept
         (libfm_parser.h:73), so a '-' raises even on a token that is then dropped */
      st.error |= kErrNegativeIndex;
    }
    if (is_label) {
      B200P_CHECK(prm, st.n_label + __popc(m_label & lt) < kRoleQueue)
      ws.q_label[st.n_label + __popc(m_label & lt)] = at;
    }
    if (is_weight) {
      B200P_CHECK(prm, st.n_weight + __popc(m_weight & lt) < kRoleQueue)
      ws.q_weight[st.n_weight + __popc(m_weight & lt)] = at;
    }
    if (FMT == kLibFM) {
      /* the field id of a triple is the token in front of its index token */
      uint32_t prev_at = __shfl_up_sync(0xffffffffu, at, 1);
      if (lane == 0) prev_at = st.carry_at;
      if (is_index) {
        const uint32_t slot = st.n_index + __popc(m_index & lt);
        B200P_CHECK(prm, slot < kRoleQueue) {
          ws.q_index[slot] = at;
          ws.q_field[slot] = prev_at;
        }
      }
    } else if (is_index) {
      const uint32_t slot = st.n_index + __popc(m_index & lt);
      B200P_CHECK(prm, slot < kRoleQueue)
      ws.q_index[slot] = at;
    }
    if (is_value) {
      B200P_CHECK(prm, st.n_value + __popc(m_value & lt) < kRoleQueue)
      ws.q_value[st.n_value + __popc(m_value & lt)] = at;
    }
    if (FMT == kLibSVM && (m_qid & me) != 0u) {
      B200P_CHECK(prm, st.n_qid + __popc(m_qid & lt) < kRoleQueue)
      ws.q_qid[st.n_qid + __popc(m_qid & lt)] = qid_at;
    }
    st.n_label += __popc(m_label);
    st.n_weight += __popc(m_weight);
    st.n_index += __popc(m_index);
    st.n_value += __popc(m_value);
    st.n_qid += __popc(m_qid);
    __syncwarp();
  }
  st.seen[kRows] += __popc(m_label);
  st.seen[kNnz] += __popc(m_index);
  st.seen[kValues] += __popc(m_value);
  st.seen[kWeights] += __popc(m_weight);
  st.seen[kQids] += __popc(m_qid);

  st.carry_pos = next_pos;
  st.carry_first = next_first;
  st.carry_at = __shfl_sync(0xffffffffu, at, static_cast<int>(count) - 1);

  if (EMIT) DrainRoleQueues<FMT, IdxT>(prm, txt, ws, st, lane, false);
}

/*
 * One group of <= 32 CSV fields, one per lane (reference csv_parser.h:73-102).  A queue entry
 * is twice the byte position of the field start, plus one when the field opens a row.
 * Every field is converted with ToFloat; column == label_column goes to label[], every other
 * column to (index = running column id without the label column, value).
 */
template <typename IdxT, bool EMIT>
__device__ __forceinline__ void FieldGroup(const WalkParams &prm, const Text &txt,
                                           WarpScratch<kCSV> &ws, WalkState<IdxT> &st, uint32_t head,
                                           uint32_t count, int lane) {
  const bool active = lane < count;
  const uint32_t entry = active ? ws.tokens[head + lane] : 0u;
  const bool row_start = active && (entry & 1u) != 0u;
  const uint32_t at = entry >> 1;
  const uint32_t rs_m = __ballot_sync(0xffffffffu, row_start);
  const uint32_t below = rs_m & LaneMaskLe();
  uint32_t col;
  if (below != 0u) {
    col = lane - (31 - __clz(below));
  } else {
    col = st.carry_col + lane + 1u;
  }
  const bool is_label = active && static_cast<int32_t>(col) == prm.label_column;
  const bool is_value = active && !is_label;
  const uint32_t lt = LaneMaskLt();
  const uint32_t m_value = __ballot_sync(0xffffffffu, is_value);
  if (EMIT) {
    const uint32_t row = st.seen[kRows] + __popc(rs_m & LaneMaskLe()) - 1u; /* row of this field */
    if (row_start) {
      B200P_CHECK(prm, st.seen[kRows] + __popc(rs_m & lt) < prm.summary->total[kRows])
      prm.offset[st.seen[kRows] + __popc(rs_m & lt)] =
          static_cast<unsigned long long>(st.seen[kNnz] + __popc(m_value & lt));
    }
    if (active) {
      const float v = ToFloatFast(txt, at);
      if (is_label) {
        B200P_CHECK(prm, row < prm.summary->total[kRows])
        prm.label[row] = v;
      } else {
        const uint32_t slot = st.seen[kNnz] + __popc(m_value & lt);
        const uint32_t idx =
            col - ((prm.label_column >= 0 && static_cast<int32_t>(col) > prm.label_column) ? 1u : 0u);
        B200P_CHECK(prm, slot < prm.summary->total[kNnz]) {
          prm.value[slot] = v;
          StoreIdx<IdxT>(prm.index, slot, idx);
        }
      }
    }
  }
  st.seen[kRows] += __popc(rs_m);
  st.seen[kNnz] += __popc(m_value);
  st.seen[kValues] += __popc(m_value);
  st.carry_col = __shfl_sync(0xffffffffu, col, static_cast<int>(count) - 1);
}

/*
 * The walk.  grid = ceil(num_ranges / kWarpsPerCta) CTAs of kCtaThreads; warp w of the grid owns
 * byte range [w * range_bytes, (w+1) * range_bytes): the lines whose first byte lies in it.
 */
template <int FMT, typename IdxT, bool EMIT, bool QID = false>
#ifdef B200P_MIN_CTAS   /* tuning experiments only; the default leaves the register budget to ptxas */
#define B200P_WALK_BOUNDS __launch_bounds__(kCtaThreads, B200P_MIN_CTAS)
#else
#define B200P_WALK_BOUNDS __launch_bounds__(kCtaThreads)
#endif
__global__ void B200P_WALK_BOUNDS WalkText(const WalkParams prm) {
  __shared__ WarpScratch<FMT> scratch[kWarpsPerCta];
  __shared__ __align__(16) uint8_t lut[256];
  for (int c = threadIdx.x; c < 256; c += kCtaThreads) lut[c] = static_cast<uint8_t>(ClassEntry<FMT>(c));
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const uint32_t range = blockIdx.x * kWarpsPerCta + warp;
  if (range >= prm.num_ranges) return;
  WarpScratch<FMT> &ws = scratch[warp];
  const Text txt = {prm.text, prm.size};

  if (EMIT) {
    /* capacity guard: if the arrays cannot hold this chunk nothing is written; the host sees
       kErrCapacity in the summary, grows the buffers and launches pass 2 again */
    const ChunkSummary *sum = prm.summary;
    if (sum->total[kRows] > prm.cap_rows || sum->total[kNnz] > prm.cap_nnz) return;
    if (range == 0 && lane == 0) prm.offset[sum->total[kRows]] = sum->total[kNnz];
  }

  const uint32_t lo = range * prm.range_bytes;
  const uint32_t hi = min(lo + prm.range_bytes, prm.size);

  /* owned bytes [begin, end): begin = first line start at or after lo, end = first line start
     at or after hi (both "one past an EOL byte"); the same cut the reference makes per OpenMP
     slice with BackFindEndLine (text_parser.h:71-77,108-116), here looking forward */
  uint32_t begin = 0;
  if (range != 0) {
    const uint32_t e = FindEol(prm.text, prm.size, lo - 1u, hi - 1u, lane);
    begin = (e == 0xffffffffu) ? hi : e + 1u;
  }

  WalkState<IdxT> st;
#pragma unroll
  for (int c = 0; c < kNumCounters; ++c) st.seen[c] = 0;
  st.carry_pos = 0;
  st.carry_first = false;
  st.carry_at = 0;
  st.carry_col = 0;
  st.n_label = st.n_weight = st.n_qid = st.n_index = st.n_value = 0;
  st.w_label = st.w_weight = st.w_qid = st.w_index = st.w_value = 0;
  st.error = 0;
  if (EMIT) {
#pragma unroll
    for (int c = 0; c < kNumCounters; ++c) {
      st.seen[c] = prm.counts[c * prm.num_ranges + range] +
                   prm.tile_base[c * prm.num_tiles + range / kScanTile];
    }
    st.w_label = st.seen[kRows];
    st.w_index = st.seen[kNnz];
    st.w_value = st.seen[kValues];
    st.w_weight = st.seen[kWeights];
    st.w_qid = st.seen[kQids];
  }

  if (begin < hi) {
    uint32_t fill = 0;          /* tokens waiting in ws.tokens */
    uint32_t end = prm.size;    /* tightened once the closing EOL is seen */
    /* classes of the byte(s) before this step's first byte.  CSV: 2 bits (separator | EOL << 1)
       of the previous byte -- the EOL before `begin` is both.  LibSVM/LibFM: class nibbles of
       the previous two bytes -- an EOL (2) in the high nibble. */
    uint32_t carry = FMT == kCSV ? 3u : 0x20u;
    bool done = false;
    uint32_t csv_fields = 0, csv_rows = 0; /* per-lane tallies of the CSV counting shortcut */
    /* interior step: wholly owned, wholly inside the chunk, and too early to hold the EOL
       that closes the region -- no masks, no bounds checks (30 of 32 steps of a 16 KiB range).
       CSV pass 2 requests its 16 bytes per lane one step ahead, so the HBM latency is spent behind
       the previous step's conversions (-3 % on B200); the other walks are tighter on registers and
       measured slower that way, so they load in place. */
    constexpr bool kAhead = (FMT == kCSV) && EMIT;
    const auto is_interior = [&](uint32_t b) { return b >= begin && b + kStepBytes <= hi - 1u; };
    const uint8_t *const my_text = prm.text + 16u * lane;
    uint4 ahead = make_uint4(0u, 0u, 0u, 0u);
    if (kAhead && is_interior(begin & ~uint32_t(kStepBytes - 1))) {
      ahead = __ldg(reinterpret_cast<const uint4 *>(my_text + (begin & ~uint32_t(kStepBytes - 1))));
    }
    for (uint32_t base = begin & ~uint32_t(kStepBytes - 1); !done; base += kStepBytes) {
      const uint32_t first_byte = base + 16u * lane;
      const bool interior = is_interior(base);
      uint4 now = ahead;
      if constexpr (kAhead) {
        if (is_interior(base + kStepBytes)) {
          ahead = __ldg(reinterpret_cast<const uint4 *>(my_text + base + kStepBytes));
        }
      } else {
        if (interior) now = __ldg(reinterpret_cast<const uint4 *>(my_text + base));
      }
      if constexpr (FMT == kCSV) {
        uint32_t cls;             /* interleaved class mask of my 16 bytes */
        uint32_t keep = kEven;    /* owned bytes, on the even bits */
        if (interior) {
          cls = ClassifyLut(now, lut, prm.four);
        } else {
          cls = ClassifyLut(Load16(prm.text, prm.size, first_byte), lut, prm.four);
          /* does the owned region close inside this step?  (first EOL at or after hi - 1) */
          const uint32_t closing = (cls >> 1) & Spread16(KeepMask(first_byte, hi - 1u, prm.size));
          const uint32_t who = __ballot_sync(0xffffffffu, closing != 0u);
          if (who != 0u) {
            const int src = __ffs(who) - 1;
            const uint32_t cs = __shfl_sync(0xffffffffu, closing, src);
            end = base + 16u * src + ((__ffs(cs) - 1) >> 1) + 1u;
            done = true;
          } else if (base + kStepBytes >= prm.size) {
            done = true;
          }
          keep = Spread16(KeepMask(first_byte, begin, end));
        }
        uint32_t up = __shfl_up_sync(0xffffffffu, cls >> 30, 1);
        if (lane == 0) up = carry;
        carry = __shfl_sync(0xffffffffu, cls >> 30, 31);
        /* a field starts after a comma or at a line start, unless that byte is an EOL
           (csv_parser.h:79-95: a trailing comma yields no extra column).  Unmasked classes are
           used for "what precedes me" so that the line start at `begin` is seen. */
        const uint32_t before = (cls << 2) | up;        /* classes of the previous byte */
        const uint32_t not_eol = ~(cls >> 1) & keep;
        const uint32_t starts = before & not_eol;
        const uint32_t row_starts = (before >> 1) & not_eol;
        const uint32_t mine = __popc(starts);
        if constexpr (!EMIT) {
          /* counting a CSV chunk needs no per-field work when the label is absent or is column 0
             (every row then owns exactly one label field): rows and fields are popcounts */
          if (prm.label_column <= 0) {
            csv_fields += mine;
            csv_rows += __popc(row_starts);
            continue;
          }
        }
        /* compact the starts, in file order, behind the tokens left from the previous step.
           Queue entry = 2 * byte position (+ 1 for a field that opens a row). */
        const uint32_t incl = WarpInclusiveSum(mine, lane);
        uint32_t slot = fill + incl - mine;
        uint32_t bits = starts;
        const uint32_t first2 = 2u * first_byte;
        while (bits != 0u) {
          const uint32_t b = __ffs(bits) - 1;     /* 2 * byte offset */
          bits &= bits - 1u;
          B200P_CHECK(prm, slot < kFieldQueue)
          ws.tokens[slot] = first2 + b + ((row_starts >> b) & 1u);
          ++slot;
        }
        fill += __shfl_sync(0xffffffffu, incl, 31);
        __syncwarp();
      } else {
        uint32_t m0, m1;                               /* class nibbles of bytes 0..7 / 8..15 */
        uint32_t k0 = 0xffffffffu, k1 = 0xffffffffu;   /* owned bytes, as nibbles */
        if (interior) {
          ClassifyLut4(now, lut, prm.sixteen, &m0, &m1);
        } else {
          ClassifyLut4(Load16(prm.text, prm.size, first_byte), lut, prm.sixteen, &m0, &m1);
          /* does the owned region close inside this step?  (first EOL at or after hi - 1) */
          const uint32_t cm = KeepMask(first_byte, hi - 1u, prm.size);
          const uint32_t c0 = (m0 >> 1) & kNibD & SpreadNibbles8(cm & 0xffu);
          const uint32_t c1 = (m1 >> 1) & kNibD & SpreadNibbles8(cm >> 8);
          uint32_t off = 0xffffffffu;                  /* byte offset of my first closing EOL */
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
          } else if (base + kStepBytes >= prm.size) {
            done = true;
          }
          const uint32_t km = KeepMask(first_byte, begin, end);
          k0 = SpreadNibbles8(km & 0xffu);
          k1 = SpreadNibbles8(km >> 8);
        }
        uint32_t up = __shfl_up_sync(0xffffffffu, m1 >> 24, 1);   /* classes of bytes -2, -1 */
        if (lane == 0) up = carry;
        carry = __shfl_sync(0xffffffffu, m1 >> 24, 31);
        /* digit-run bytes I own; a run cannot cross `begin` or `end` (both follow an EOL byte) */
        const uint32_t d0 = m0 & kNibD & k0, d1 = m1 & kNibD & k1;
        const uint32_t s0 = d0 & ~((d0 << 4) | ((up >> 4) & 1u));
        const uint32_t s1 = d1 & ~((d1 << 4) | (d0 >> 28));
        const uint32_t mine = __popc(s0) + __popc(s1);
        /* compact the run starts.  Queue entry = 4 * byte position + separator kind, the kind read
           off the (unmasked) classes of the two bytes in front of the token. */
        const uint32_t incl = WarpInclusiveSum(mine, lane);
        uint32_t slot = fill + incl - mine;
        const uint32_t p_lo = up | (m0 << 8), p_mid = (m0 >> 24) | (m1 << 8), p_hi = m1 >> 24;
        const uint32_t first4 = 4u * first_byte;
        uint32_t bits = s0;
        while (bits != 0u) {
          const uint32_t b = __ffs(bits) - 1;     /* 4 * byte offset */
          bits &= bits - 1u;
          B200P_CHECK(prm, slot < kTokenQueue)
          ws.tokens[slot] = first4 + b + SeparatorKind(__funnelshift_r(p_lo, p_mid, b) & 0xffu);
          ++slot;
        }
        bits = s1;
        while (bits != 0u) {
          const uint32_t b = __ffs(bits) - 1;
          bits &= bits - 1u;
          B200P_CHECK(prm, slot < kTokenQueue)
          ws.tokens[slot] = first4 + 32u + b + SeparatorKind(__funnelshift_r(p_mid, p_hi, b) & 0xffu);
          ++slot;
        }
        fill += __shfl_sync(0xffffffffu, incl, 31);
        __syncwarp();
      }
      /* handle full groups (all of them when the region is finished) */
      uint32_t head = 0;
      while (fill - head >= 32u || (done && fill > head)) {
        const uint32_t count = min(fill - head, 32u);
        if constexpr (FMT == kCSV) {
          FieldGroup<IdxT, EMIT>(prm, txt, ws, st, head, count, lane);
        } else {
          TokenGroup<FMT, IdxT, EMIT, QID>(prm, txt, ws, st, head, count, lane);
        }
        head += count;
      }
      /* move the < 32 left-over tokens to the front */
      if (head != 0u) {
        const uint32_t rest = fill - head;
        uint32_t moved = 0;
        if (lane < rest) moved = ws.tokens[head + lane];
        __syncwarp();
        if (lane < rest) ws.tokens[lane] = moved;
        __syncwarp();
        fill = rest;
      }
    }
    if constexpr (EMIT && FMT != kCSV) DrainRoleQueues<FMT, IdxT>(prm, txt, ws, st, lane, true);
    if constexpr (FMT == kCSV && !EMIT) {
      if (prm.label_column <= 0) {
        const uint32_t rows = __reduce_add_sync(0xffffffffu, csv_rows);
        const uint32_t fields = __reduce_add_sync(0xffffffffu, csv_fields);
        st.seen[kRows] = rows;
        st.seen[kNnz] = fields - (prm.label_column == 0 ? rows : 0u);
        st.seen[kValues] = st.seen[kNnz];
      }
    }
  }

#ifdef B200P_CHECKED
  if (EMIT) {
    /* pass 2 must have produced exactly what pass 1 counted for this range */
    bool same = true;
#pragma unroll
    for (int c = 0; c < kNumCounters; ++c) {
      const unsigned long long next =
          range + 1 < prm.num_ranges
              ? static_cast<unsigned long long>(prm.counts[c * prm.num_ranges + range + 1]) +
                    prm.tile_base[c * prm.num_tiles + (range + 1) / kScanTile]
              : prm.summary->total[c];
      same = same && (static_cast<unsigned long long>(st.seen[c]) == next);
    }
    if (FMT != kCSV) { /* the role queues must have drained to the same positions */
      same = same && st.w_label == st.seen[kRows] && st.w_index == st.seen[kNnz] &&
             st.w_value == st.seen[kValues] && st.w_weight == st.seen[kWeights] &&
             st.w_qid == st.seen[kQids];
    }
    if (!same) st.error |= kErrInternal;
  }
#endif
  if (EMIT) {
    const uint32_t err = __reduce_or_sync(0xffffffffu, st.error);
    if (err != 0u && lane == 0) atomicOr(&prm.summary->error, err);
  } else {
    if (lane < kNumCounters) {
      uint32_t v = 0;
#pragma unroll
      for (int c = 0; c < kNumCounters; ++c) v = (lane == c) ? st.seen[c] : v;
      prm.counts[lane * prm.num_ranges + range] = v;
    }
  }
}

/*
 * Exclusive prefix sums of the per-range counters, in two small launches.
 *
 * ScanTiles: CTA t scans ranges [t*1024, (t+1)*1024) of all five counters in place (coalesced:
 * thread i <-> range t*1024+i) and stores the tile's totals.  ScanFinish: one CTA turns the tile
 * totals into exclusive tile bases (in place), the chunk totals and the RowBlock consistency
 * check of the reference (row_block.h:177).  Pass 2 adds counts[range] + tile_base[range/1024].
 */
__device__ __forceinline__ uint32_t BlockExclusiveScan(uint32_t v, uint32_t *warp_sums /*[32]*/,
                                                       uint32_t *total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t incl = WarpInclusiveSum(v, lane);
  if (lane == 31) warp_sums[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    const uint32_t w = warp_sums[lane];
    const uint32_t wi = WarpInclusiveSum(w, lane);
    warp_sums[lane] = wi - w; /* exclusive base of each warp */
    if (lane == 31) *total = wi;
  }
  __syncthreads();
  const uint32_t r = incl - v + warp_sums[warp];
  return r;
}

__global__ void __launch_bounds__(kScanTile) ScanTiles(uint32_t *counts, uint32_t num_ranges,
                                                       uint32_t *tile_base, uint32_t num_tiles) {
  __shared__ uint32_t warp_sums[kNumCounters][32];
  __shared__ uint32_t totals[kNumCounters];
  const uint32_t i = blockIdx.x * kScanTile + threadIdx.x;
#pragma unroll
  for (int c = 0; c < kNumCounters; ++c) {
    const uint32_t v = i < num_ranges ? counts[c * num_ranges + i] : 0u;
    const uint32_t ex = BlockExclusiveScan(v, warp_sums[c], &totals[c]);
    if (i < num_ranges) counts[c * num_ranges + i] = ex;
  }
  __syncthreads();
  if (threadIdx.x < kNumCounters) tile_base[threadIdx.x * num_tiles + blockIdx.x] = totals[threadIdx.x];
}

__global__ void __launch_bounds__(kScanTile) ScanFinish(uint32_t *tile_base, uint32_t num_tiles,
                                                        uint32_t num_ranges, ChunkSummary *summary,
                                                        const uint32_t *extra_error) {
  __shared__ uint32_t warp_sums[32];
  __shared__ uint32_t total;
  __shared__ unsigned long long grand[kNumCounters];
  for (int c = 0; c < kNumCounters; ++c) {
    uint32_t carry = 0;
    for (uint32_t base = 0; base < num_tiles; base += kScanTile) {
      const uint32_t i = base + threadIdx.x;
      const uint32_t v = i < num_tiles ? tile_base[c * num_tiles + i] : 0u;
      const uint32_t ex = BlockExclusiveScan(v, warp_sums, &total);
      if (i < num_tiles) tile_base[c * num_tiles + i] = carry + ex;
      carry += total;
      __syncthreads();
    }
    if (threadIdx.x == 0) grand[c] = carry;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int err = 0;
#pragma unroll
    for (int c = 0; c < kNumCounters; ++c) summary->total[c] = grand[c];
    if (grand[kValues] != 0ull && grand[kValues] != grand[kNnz]) err |= kErrValueCount;
    if (extra_error != NULL) err |= *extra_error;   /* bits the collecting kernel of split_kernels.cuh raised */
    summary->error = err;
    summary->ranges = num_ranges;
  }
}

/*
 * Largest element of an index / field array, on demand (b200p_block_max): what
 * RowBlockContainer::Push accumulates per element on the CPU (reference row_block.h:143,153) and
 * RowBlockIter::NumCol() reports.  Kept out of the emit kernel: streaming Parser users never ask,
 * and the array is still hot in L2 when a RowBlockIter back-end does.
 */
template <typename IdxT>
__global__ void __launch_bounds__(256) MaxOfArray(const IdxT *values, unsigned long long n,
                                                  unsigned long long *result) {
  IdxT m = 0;
  for (unsigned long long i = blockIdx.x * 256ull + threadIdx.x; i < n; i += gridDim.x * 256ull) {
    m = MaxOf<IdxT>(m, values[i]);
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    m = MaxOf<IdxT>(m, static_cast<IdxT>(__shfl_xor_sync(0xffffffffu, m, d)));
  }
  if ((threadIdx.x & 31) == 0 && m != 0) {
    const volatile unsigned long long *seen = result;
    if (static_cast<unsigned long long>(m) > *seen) atomicMax(result, static_cast<unsigned long long>(m));
  }
}

/*
 * LibFM field ids for the trip over PCIe: one byte each behind a 256-byte header whose first
 * word is set when some id does not fit (b200p_next then copies the wide array after all).
 * Field ids are small by nature (the reference's own test data has 39); as IndexType they are 4
 * or 8 bytes of every entry sent back to the host.
 */
constexpr unsigned kNarrowHeader = 256;
template <typename IdxT>
__global__ void __launch_bounds__(256) NarrowFields(const IdxT *field, unsigned long long n,
                                                    unsigned char *narrow) {
  bool wide = false;
  for (unsigned long long i = blockIdx.x * 256ull + threadIdx.x; i < n; i += gridDim.x * 256ull) {
    const IdxT v = field[i];
    narrow[kNarrowHeader + i] = static_cast<unsigned char>(v);
    wide = wide || v > static_cast<IdxT>(255);
  }
  if (__any_sync(0xffffffffu, wide) && (threadIdx.x & 31) == 0) {
    atomicOr(reinterpret_cast<unsigned int *>(narrow), 1u);
  }
}

}  // namespace b200p
#endif  // DMLC_CORE_B200_TEXT_KERNELS_CUH_
