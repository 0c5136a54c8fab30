/*
 * CollectRanges + ConvertRanges: the front-end runs once, its result goes through memory.
 *
 * The two-pass walk (text_kernels.cuh) classifies, compacts and pairs the tokens of a chunk twice;
 * the one-launch kernel (tile_kernels.cuh) does it once but pays for that with waits: a tile's
 * warps stand at a barrier until the look-back has the tile's position, and the look-back in turn
 * waits for the slowest of the dozens of tiles in flight in front of it.  Here the two halves are
 * separate launches with nothing to wait for inside either of them:
 *
 *   CollectRanges   warp w walks byte range w of the chunk (text straight from global memory, 16
 *                   bytes per lane and step, like pass 1 of the walk), pairs tokens into roles
 *                   (TokenRoles) and writes each token's position -- 16 bits, relative to the
 *                   range -- into the queue of its role, in a per-range block of global scratch;
 *                   the five counts go where ScanTiles expects them
 *   ScanTiles + ScanFinish (text_kernels.cuh)   exclusive prefixes, totals, consistency check
 *   ConvertRanges   warp w reads its queues back and converts, one role at a time, 32 tokens per
 *                   trip: every lane runs the same converter, every store is 32 consecutive
 *                   elements of one CSR array
 *
 * The queues cost DRAM traffic (2 bytes per token each way, about half the text size together)
 * on a path that sits far below the HBM roofline because it is bound by instruction issue; they
 * buy a front-end that runs once and two kernels without a single barrier.
 * A range whose lines span more than 64 KiB (LibSVM / LibFM), or which holds more tokens or rows
 * than its queues, raises kErrFallback and the chunk goes through the two-pass kernels; so does
 * CSV with the label in a later column than the first (it needs a column id per field).
 */
#ifndef DMLC_CORE_B200_SPLIT_KERNELS_CUH_
#define DMLC_CORE_B200_SPLIT_KERNELS_CUH_

#include "text_kernels.cuh"

namespace b200p {

/* checked build (-DB200P_CHECKED): every queue write and every CSR store of these kernels is
   bounds-checked; a violation sets kErrInternal instead of writing (see text_kernels.cuh) */
#ifdef B200P_CHECKED
#define B200P_SPLIT_CHECK(sp, cond) \
  if (!(cond)) { atomicOr((sp).give_up, static_cast<uint32_t>(kErrInternal)); } else
#define B200P_OUT_CHECK(prm, cond) \
  if (!(cond)) { atomicOr(&(prm).summary->error, static_cast<unsigned int>(kErrInternal)); } else
#else
#define B200P_SPLIT_CHECK(sp, cond)
#define B200P_OUT_CHECK(prm, cond)
#endif

struct SplitParams {
  WalkParams walk;          /* text, ranges, counts / tile_base / summary, outputs: as for the walk */
  uint16_t *queues;         /* num_ranges blocks of queue_stride entries */
  uint32_t queue_stride;    /* entries per range block */
  uint32_t cap_pool;        /* index entries from the front, value entries from the back */
  uint32_t cap_rows;
  uint32_t *give_up;        /* set to kErrFallback (| kErrDense) when a range does not fit */
};

/* entries one range block holds, and where the queues sit in it */
template <int FMT>
__host__ __device__ inline uint32_t SplitQueueEntries(uint32_t cap_pool, uint32_t cap_rows) {
  uint32_t n = cap_pool + 3u * cap_rows;                 /* pool, row_at, row_nnz, weight */
  if (FMT == kLibSVM) n += cap_rows;                     /* qid */
  if (FMT == kLibFM) n += cap_pool / 2u;                 /* field */
  return (n + 7u) & ~7u;
}
struct SplitOffsets {
  uint32_t row_at, row_nnz, weight, qid, field;
};
template <int FMT>
__device__ __forceinline__ SplitOffsets SplitLayout(uint32_t cap_pool, uint32_t cap_rows) {
  SplitOffsets o;
  o.row_at = cap_pool;
  o.row_nnz = o.row_at + cap_rows;
  o.weight = o.row_nnz + cap_rows;
  o.qid = o.weight + cap_rows;
  o.field = FMT == kLibSVM ? o.qid + cap_rows : o.qid;
  return o;
}

/* what ConvertRanges needs to know about a range besides the five counts of the scan */
#ifndef B200P_SPLIT_WARPS
#define B200P_SPLIT_WARPS 8
#endif
constexpr int kSplitWarps = B200P_SPLIT_WARPS;
constexpr int kSplitThreads = kSplitWarps * 32;

/* resident CTAs per SM the register budget of CollectRanges is held to.  Measured on B200 (500 MB
   per shape, ms per launch, LibSVM / LibFM u64 / Criteo): 4 CTAs (64 registers) 0.586 / 0.666 /
   0.617, 5 CTAs (48) 0.642 / 0.710 / 0.664, 6 CTAs (40) 0.662 / 0.775 / 0.686 -- the spills cost
   more instructions than the extra warps hide; 3 CTAs (80 registers, no value rebuilt) are 8 %
   slower than 4: the kernel wants its 32 warps. */
#ifndef B200P_COLLECT_CTAS
#define B200P_COLLECT_CTAS (32 / B200P_SPLIT_WARPS)
#endif
template <int FMT, typename IdxT, bool QID>
__global__ void __launch_bounds__(kSplitThreads, B200P_COLLECT_CTAS) CollectRanges(const SplitParams sp) {
  __shared__ uint32_t rings[kSplitWarps][kTokenQueue];
  __shared__ __align__(16) uint8_t lut[256];
  /* separator kind by the class nibbles of the two bytes in front of a token: one LDS per token
     where SeparatorKind() costs seven instructions, five of them on the ALU pipe that bounds
     this kernel */
  __shared__ __align__(16) uint8_t kind_of[256];
  const WalkParams &prm = sp.walk;
  for (int c = threadIdx.x; c < 256; c += kSplitThreads) {
    lut[c] = static_cast<uint8_t>(ClassEntry<FMT>(c));
    kind_of[c] = static_cast<uint8_t>(SeparatorKind(c));
  }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const uint32_t range = blockIdx.x * kSplitWarps + warp;
  if (range >= prm.num_ranges) return;
  uint32_t *const tokens = rings[warp];
  const Text txt = {prm.text, prm.size};
  uint16_t *q16 = sp.queues + static_cast<size_t>(range) * sp.queue_stride;
  const SplitOffsets at = SplitLayout<FMT>(sp.cap_pool, sp.cap_rows);

  uint32_t lo = range * prm.range_bytes;
  /* held in registers: left to itself the compiler rebuilds the block's address (a 64-bit
     multiply-add and two adds) and `lo` from the kernel parameters at every store of the loop */
  asm volatile("" : "+l"(q16), "+r"(lo));
  const uint32_t hi = min(lo + prm.range_bytes, prm.size);
  uint32_t n_rows = 0, n_index = 0, n_value = 0, n_weight = 0, n_qid = 0;
  uint32_t error = 0;
  bool give_up = false, dense = false;

  /* owned bytes [begin, end): the lines that start in [lo, hi) (text_parser.h:104-118) */
  uint32_t begin = 0;
  if (range != 0) {
    const uint32_t e = FindEol(prm.text, prm.size, lo - 1u, hi - 1u, lane);
    begin = (e == 0xffffffffu) ? hi : e + 1u;
  }
  if (begin < hi) {
    uint32_t fill = 0;
    uint32_t end = prm.size;
    uint32_t carry = 0x20u;        /* class nibbles of the two bytes in front: an EOL before `begin` */
    int carry_pos = 0;
    bool carry_first = false;
    uint32_t carry_at = 0;
    bool done = false;
    const uint32_t lt = LaneMaskLt();
    const uint32_t me = 1u << lane;
    const uint8_t *const my_text = prm.text + 16u * lane;
    uint32_t v_end = sp.cap_pool - 1u;    /* value entries grow down from the end of the pool */

    /* the text of an interior step is requested one step ahead, so that the trip to HBM is spent
       behind the previous step's token work */
    const auto is_interior = [&](uint32_t b) { return b >= begin && b + kStepBytes <= hi - 1u; };
    uint4 ahead = make_uint4(0u, 0u, 0u, 0u);
    if (is_interior(begin & ~uint32_t(kStepBytes - 1))) {
      ahead = __ldg(reinterpret_cast<const uint4 *>(my_text + (begin & ~uint32_t(kStepBytes - 1))));
    }
    for (uint32_t base = begin & ~uint32_t(kStepBytes - 1); !done; base += kStepBytes) {
      const uint32_t first_byte = base + 16u * lane;
      const bool interior = is_interior(base);
      const uint4 now = ahead;
      if (is_interior(base + kStepBytes)) ahead = __ldg(reinterpret_cast<const uint4 *>(my_text + base + kStepBytes));
      uint32_t m0, m1;
      uint32_t k0 = 0xffffffffu, k1 = 0xffffffffu;
      if (interior) {
        ClassifyLut4(now, lut, prm.sixteen, &m0, &m1);
      } else {
        if (base + kStepBytes - lo > 0xffffu) {
          /* positions are stored as 16-bit offsets from `lo`: a line that carries the range past
             64 KiB is left to the two-pass kernels (a range is at most 32 KiB, so only a step
             behind its last byte -- never an interior one -- can get there) */
          give_up = true;
          break;
        }
        ClassifyLut4(Load16(prm.text, prm.size, first_byte), lut, prm.sixteen, &m0, &m1);
        const uint32_t cm = KeepMask(first_byte, hi - 1u, prm.size);
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
        } else if (base + kStepBytes >= prm.size) {
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
      /* rolled: unrolled sevenfold (the compiler's choice) the two loops run 1-3 % slower */
      uint32_t bits = s0;
#pragma unroll 1
      while (bits != 0u) {
        const uint32_t b = __ffs(bits) - 1;
        bits &= bits - 1u;
        B200P_SPLIT_CHECK(sp, slot < kTokenQueue)
        tokens[slot] = first4 + b + kind_of[__funnelshift_r(p_lo, p_mid, b) & 0xffu];
        ++slot;
      }
      bits = s1;
#pragma unroll 1
      while (bits != 0u) {
        const uint32_t b = __ffs(bits) - 1;
        bits &= bits - 1u;
        B200P_SPLIT_CHECK(sp, slot < kTokenQueue)
        tokens[slot] = first4 + 32u + b + kind_of[__funnelshift_r(p_mid, p_hi, b) & 0xffu];
        ++slot;
      }
      fill += __shfl_sync(0xffffffffu, incl, 31);
      __syncwarp();

      /* room for whatever the queued tokens turn out to be: each of them adds at most one entry
         to the pool, to the row queues, to the field queue */
      if (n_index + n_value + fill > sp.cap_pool || n_rows + fill > sp.cap_rows ||
          (FMT == kLibFM && n_index + fill > sp.cap_pool / 2u)) {
        give_up = true;
        dense = true;
        break;
      }
      /* full groups of 32 tokens (all of them once the owned lines are finished) */
      uint32_t head = 0;
      const uint32_t groups = done ? (fill + 31u) >> 5 : fill >> 5;
      for (uint32_t g = 0; g < groups; ++g) {
        const uint32_t count = min(fill - head, 32u);
        const GroupRoles r = TokenRoles<FMT, QID>(txt, tokens, head, count, lane, carry_pos, carry_first);
        /* one 16-bit store per token, into the queue of its role; the queues' running ends are
           kept as offsets into the range's block */
        const uint16_t rel = static_cast<uint16_t>(r.at - lo);
        const uint32_t k_index = n_index + __popc(r.m_index & lt);
        const uint32_t k_value = v_end - __popc(r.m_value & lt);
        const bool is_index = (r.m_index & me) != 0u, is_value = (r.m_value & me) != 0u;
        if (is_index || is_value) {
          B200P_SPLIT_CHECK(sp, (is_index ? k_index : k_value) < sp.cap_pool && k_index <= k_value + 1u)
          q16[is_index ? k_index : k_value] = rel;    /* a token is never both */
        }
        /* label, weight, qid: one test for the groups that hold none of them */
        if ((r.m_label | r.m_weight | (FMT == kLibSVM ? r.m_qid : 0u)) != 0u) {
          if (r.m_label != 0u) {           /* a row starts in this group */
            if ((r.m_label & me) != 0u) {
              const uint32_t k = n_rows + __popc(r.m_label & lt);
              B200P_SPLIT_CHECK(sp, k < sp.cap_rows) {
                q16[at.row_at + k] = rel;
                q16[at.row_nnz + k] = static_cast<uint16_t>(n_index + __popc(r.m_index & lt));
              }
            }
            n_rows += __popc(r.m_label);
          }
          if (r.m_weight != 0u) {
            if ((r.m_weight & me) != 0u) q16[at.weight + n_weight + __popc(r.m_weight & lt)] = rel;
            n_weight += __popc(r.m_weight);
          }
          if (FMT == kLibSVM && r.m_qid != 0u) {
            if ((r.m_qid & me) != 0u) q16[at.qid + n_qid + __popc(r.m_qid & lt)] = static_cast<uint16_t>(r.qid_at - lo);
            n_qid += __popc(r.m_qid);
          }
        }
        if (FMT == kLibFM) {
          /* a '-' on a field id raises even when the triple is dropped (libfm_parser.h:73);
             kept ids report it when they are converted, dropped ones are looked at here */
          const uint32_t maybe_dropped = r.m_field & ~(r.m_index >> 1);
          if (maybe_dropped != 0u) {
            if ((maybe_dropped & me) != 0u && txt.at(r.at) == '-') error |= kErrNegativeIndex;
          }
          /* the field id of an entry is the token in front of its index token */
          uint32_t prev_at = __shfl_up_sync(0xffffffffu, r.at, 1);
          if (lane == 0) prev_at = carry_at;
          if (is_index) {
            B200P_SPLIT_CHECK(sp, k_index < sp.cap_pool / 2u && at.field + k_index < sp.queue_stride)
            q16[at.field + k_index] = static_cast<uint16_t>(prev_at - lo);
          }
          carry_at = __shfl_sync(0xffffffffu, r.at, static_cast<int>(count) - 1);
        }
        n_index += __popc(r.m_index);
        const uint32_t nv = __popc(r.m_value);
        n_value += nv;
        v_end -= nv;
        carry_pos = r.next_pos;
        carry_first = r.next_first;
        head += count;
      }
      if (head != 0u) {
        const uint32_t rest = fill - head;
        uint32_t moved = 0;
        if (lane < rest) moved = tokens[head + lane];
        __syncwarp();
        if (lane < rest) tokens[lane] = moved;
        __syncwarp();
        fill = rest;
      }
    }
  }
  if (lane < kNumCounters) {
    uint32_t v = n_rows;
    v = lane == kNnz ? n_index : v;
    v = lane == kValues ? n_value : v;
    v = lane == kWeights ? n_weight : v;
    v = lane == kQids ? n_qid : v;
    prm.counts[lane * prm.num_ranges + range] = give_up ? 0u : v;
  }
  error = __reduce_or_sync(0xffffffffu, error);
  if (give_up) error |= kErrFallback | (dense ? static_cast<uint32_t>(kErrDense) : 0u);
  if (error != 0u && lane == 0) atomicOr(sp.give_up, error);
}

/* the queues of CollectRanges drained into the CSR arrays */
/* resident CTAs per SM the register budget of the convert kernels is held to: 6 (40 registers,
   48 warps) against the 46 registers / 5 CTAs the u64 instantiations take when left alone is worth
   1-2 % on LibSVM; 8 (32 registers, spills) costs 7-14 % */
#ifndef B200P_CONVERT_CTAS
#define B200P_CONVERT_CTAS (48 / B200P_SPLIT_WARPS)
#endif
template <int FMT, typename IdxT>
__global__ void __launch_bounds__(kSplitThreads, B200P_CONVERT_CTAS) ConvertRanges(const SplitParams sp) {
  const WalkParams &prm = sp.walk;
  const int lane = threadIdx.x & 31;
  const uint32_t range = blockIdx.x * kSplitWarps + (threadIdx.x >> 5);
  if (range >= prm.num_ranges) return;
  const ChunkSummary *sum = prm.summary;
  /* nothing is written when the chunk goes to the two-pass kernels or does not fit the arrays */
  if ((sum->error & kErrFallback) != 0u) return;
  if (sum->total[kRows] > prm.cap_rows || sum->total[kNnz] > prm.cap_nnz) return;
  if (range == 0 && lane == 0) prm.offset[sum->total[kRows]] = sum->total[kNnz];

  uint32_t b[kNumCounters], n[kNumCounters];
#pragma unroll
  for (int c = 0; c < kNumCounters; ++c) {
    b[c] = prm.counts[c * prm.num_ranges + range] + prm.tile_base[c * prm.num_tiles + range / kScanTile];
    const unsigned long long next =
        range + 1 < prm.num_ranges
            ? static_cast<unsigned long long>(prm.counts[c * prm.num_ranges + range + 1]) +
                  prm.tile_base[c * prm.num_tiles + (range + 1) / kScanTile]
            : sum->total[c];
    n[c] = static_cast<uint32_t>(next - b[c]);
  }
  const Text txt = {prm.text, prm.size};
  const uint32_t lo = range * prm.range_bytes;
  const uint16_t *q16 = sp.queues + static_cast<size_t>(range) * sp.queue_stride;
  asm volatile("" : "+l"(q16));     /* kept in registers, see CollectRanges */
  const SplitOffsets at = SplitLayout<FMT>(sp.cap_pool, sp.cap_rows);
  IdxT *const index = reinterpret_cast<IdxT *>(prm.index);

  bool neg = false;
  B200P_OUT_CHECK(prm, b[kRows] + n[kRows] <= sum->total[kRows] && b[kNnz] + n[kNnz] <= sum->total[kNnz] &&
                           b[kValues] + n[kValues] <= sum->total[kValues] && b[kWeights] + n[kWeights] <= sum->total[kWeights] &&
                           b[kQids] + n[kQids] <= sum->total[kQids] && n[kNnz] + n[kValues] <= sp.cap_pool &&
                           n[kRows] <= sp.cap_rows) {
  for (uint32_t i = lane; i < n[kRows]; i += 32u) {
    prm.label[b[kRows] + i] = ToFloatFast(txt, lo + q16[at.row_at + i]);
    prm.offset[b[kRows] + i] = static_cast<unsigned long long>(b[kNnz] + q16[at.row_nnz + i]);
  }
  for (uint32_t i = lane; i < n[kWeights]; i += 32u) {
    prm.weight[b[kWeights] + i] = ToFloatFast(txt, lo + q16[at.weight + i]);
  }
  if constexpr (FMT == kLibSVM) {
    for (uint32_t i = lane; i < n[kQids]; i += 32u) {
      prm.qid[b[kQids] + i] = ToQid(txt, lo + q16[at.qid + i]);
    }
  }
  /* Entries and their values in one sweep: the value of entry i stands right behind its index in
     the text, so the lines one converter pulled into L1 are the ones the next one wants.  The
     queue entries of the next trip are read one trip ahead and the text they point at is
     requested from L2 (prefetch), so that neither the queue read nor the token windows of a trip
     wait for memory. */
  const uint32_t sweep = max(n[kNnz], n[kValues]);
  const uint8_t *const text_lo = prm.text + lo;
  const uint32_t last_rel = prm.size - 1u - lo;       /* prefetches stay inside the chunk */
  const auto prefetch = [&](uint32_t rel) {
    asm volatile("prefetch.global.L1 [%0];" ::"l"(text_lo + rel));
    asm volatile("prefetch.global.L1 [%0];" ::"l"(text_lo + min(rel + 20u, last_rel)));
  };
  /* pipeline: positions of trip t+1 (a*) and t+2 (b*) are in registers, those of t+3 are being
     read; the text of trip t+2 is requested while trip t is converted */
  const auto index_at = [&](uint32_t k) -> uint32_t { return k < n[kNnz] ? q16[k] : 0u; };
  const auto field_at = [&](uint32_t k) -> uint32_t { return (FMT == kLibFM && k < n[kNnz]) ? q16[at.field + k] : 0u; };
  const auto value_at = [&](uint32_t k) -> uint32_t { return k < n[kValues] ? q16[sp.cap_pool - 1u - k] : 0u; };
  uint32_t p_index = index_at(lane), p_field = field_at(lane), p_value = value_at(lane);
  uint32_t a_index = index_at(lane + 32u), a_field = field_at(lane + 32u), a_value = value_at(lane + 32u);
  uint32_t b_index = index_at(lane + 64u), b_field = field_at(lane + 64u), b_value = value_at(lane + 64u);
  for (uint32_t i = lane; i < sweep; i += 32u) {
    const uint32_t c_index = index_at(i + 96u), c_field = field_at(i + 96u), c_value = value_at(i + 96u);
    if (i + 64u < n[kNnz]) prefetch(FMT == kLibFM ? b_field : b_index);
    if (i + 64u < n[kValues]) prefetch(b_value);
    if (i < n[kNnz]) {
      if constexpr (FMT == kLibFM) {
        reinterpret_cast<IdxT *>(prm.field)[b[kNnz] + i] =
            static_cast<IdxT>(ToUintFast<sizeof(IdxT) == 8>(txt, lo + p_field, &neg));
      }
      index[b[kNnz] + i] = static_cast<IdxT>(ToUintFast<sizeof(IdxT) == 8>(txt, lo + p_index, &neg));
    }
    if (i < n[kValues]) prm.value[b[kValues] + i] = ToFloatFast(txt, lo + p_value);
    p_index = a_index; p_field = a_field; p_value = a_value;
    a_index = b_index; a_field = b_field; a_value = b_value;
    b_index = c_index; b_field = c_field; b_value = c_value;
  }
  }   /* B200P_OUT_CHECK */
  if (__any_sync(0xffffffffu, neg) && lane == 0) atomicOr(&prm.summary->error, static_cast<unsigned int>(kErrNegativeIndex));
}

/* ---------------------------------------------------------------- CSV ------------------------- */
/*
 * CSV (label in column 0, or no label column): rows need no pairing, so a warp owns the FIELDS
 * that start in its range, whatever row they belong to (reference csv_parser.h:73-102: every
 * field is a float; column label_column is the label, every other column an (index, value)
 * entry whose index is its position among the row's value columns).  A range block holds the
 * value positions, the first field of every row that starts in the range, the number of values
 * in front of each of those rows, and -- for the column ids of values that continue a row begun
 * in an earlier range -- how many values follow the range's last row start.
 */
__host__ __device__ inline uint32_t SplitCsvEntries(uint32_t cap_pool, uint32_t cap_rows) {
  return (cap_pool + 2u * cap_rows + 2u + 7u) & ~7u;    /* values, row_at, row_nnz, 2 entries of meta */
}

template <typename IdxT>
__global__ void __launch_bounds__(kSplitThreads) CollectFieldRanges(const SplitParams sp) {
  __shared__ __align__(16) uint8_t lut[256];
  const WalkParams &prm = sp.walk;
  for (int c = threadIdx.x; c < 256; c += kSplitThreads) lut[c] = static_cast<uint8_t>(ClassEntry<kCSV>(c));
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const uint32_t range = blockIdx.x * kSplitWarps + (threadIdx.x >> 5);
  if (range >= prm.num_ranges) return;
  uint16_t *q16 = sp.queues + static_cast<size_t>(range) * sp.queue_stride;
  asm volatile("" : "+l"(q16));     /* kept in registers, see CollectRanges */
  const uint32_t o_row_at = sp.cap_pool, o_row_nnz = sp.cap_pool + sp.cap_rows, o_meta = sp.cap_pool + 2u * sp.cap_rows;
  const uint32_t lo = range * prm.range_bytes;
  const uint32_t hi = min(lo + prm.range_bytes, prm.size);
  const bool labelled = prm.label_column == 0;
  uint32_t n_rows = 0, n_value = 0, last_row_nnz = 0;
  bool has_row = false, give_up = false;
  /* class of the byte in front of the range (separator | EOL << 1); the chunk starts a line */
  uint32_t carry = lo == 0u ? 3u : lut[__ldg(prm.text + lo - 1u)];
  const uint8_t *const my_text = prm.text + 16u * lane;
  const auto whole = [&](uint32_t b) { return b + kStepBytes <= hi; };   /* hi <= size: the loads stay inside the chunk */
  uint4 ahead = make_uint4(0u, 0u, 0u, 0u);
  if (whole(lo)) ahead = __ldg(reinterpret_cast<const uint4 *>(my_text + lo));
  for (uint32_t base = lo; base < hi; base += kStepBytes) {
    const uint32_t first_byte = base + 16u * lane;
    uint4 now = ahead;
    if (!whole(base)) now = Load16(prm.text, prm.size, first_byte);
    if (whole(base + kStepBytes)) ahead = __ldg(reinterpret_cast<const uint4 *>(my_text + base + kStepBytes));
    const uint32_t cls = ClassifyLut(now, lut, prm.four);
    uint32_t keep = kEven;
    if (!whole(base)) keep = Spread16(KeepMask(first_byte, lo, hi));
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
    const uint32_t step_both = __shfl_sync(0xffffffffu, both, 31);
    const uint32_t step_values = step_both & 0xffffu, step_rows = step_both >> 16;
    if (n_rows + step_rows > sp.cap_rows) {    /* cap_pool = range / 2 always holds the fields */
      give_up = true;
      break;
    }
    const uint32_t vbase = n_value + (both & 0xffffu) - nv;
    const uint32_t rel2 = 2u * (first_byte - lo);
    uint32_t vslot = vbase;
    uint32_t bits = values;
    while (bits != 0u) {
      const uint32_t b = __ffs(bits) - 1;          /* 2 * byte offset */
      bits &= bits - 1u;
      q16[vslot] = static_cast<uint16_t>((rel2 + b) >> 1);
      ++vslot;
    }
    if (step_rows != 0u) {
      uint32_t rslot = n_rows + (both >> 16) - nr;
      bits = row_starts;
      while (bits != 0u) {
        const uint32_t b = __ffs(bits) - 1;
        bits &= bits - 1u;
        q16[o_row_at + rslot] = static_cast<uint16_t>((rel2 + b) >> 1);
        q16[o_row_nnz + rslot] = static_cast<uint16_t>(vbase + __popc(values & ((1u << b) - 1u)));
        ++rslot;
      }
      const uint32_t who = __ballot_sync(0xffffffffu, nr != 0u);
      const int src = 31 - __clz(who);
      const uint32_t top = 31u - __clz(row_starts | 1u);
      last_row_nnz = __shfl_sync(0xffffffffu, vbase + __popc(values & ((1u << top) - 1u)), src);
      has_row = true;
    }
    n_value += step_values;
    n_rows += step_rows;
  }
  if (give_up) {
    n_rows = n_value = 0u;
    if (lane == 0) atomicOr(sp.give_up, static_cast<uint32_t>(kErrFallback | kErrDense));
  }
  if (lane == 0) {
    /* values behind the range's last row start (all of them when no row starts here) */
    const uint32_t tail = has_row ? n_value - last_row_nnz : n_value;
    q16[o_meta] = static_cast<uint16_t>(tail);
    q16[o_meta + 1u] = has_row ? 1u : 0u;
  }
  if (lane < kNumCounters) {
    const uint32_t v = lane == kRows ? n_rows : ((lane == kNnz || lane == kValues) ? n_value : 0u);
    prm.counts[lane * prm.num_ranges + range] = v;
  }
}

template <typename IdxT>
__global__ void __launch_bounds__(kSplitThreads, B200P_CONVERT_CTAS) ConvertFieldRanges(const SplitParams sp) {
  const WalkParams &prm = sp.walk;
  const int lane = threadIdx.x & 31;
  const uint32_t range = blockIdx.x * kSplitWarps + (threadIdx.x >> 5);
  if (range >= prm.num_ranges) return;
  const ChunkSummary *sum = prm.summary;
  if ((sum->error & kErrFallback) != 0u) return;
  if (sum->total[kRows] > prm.cap_rows || sum->total[kNnz] > prm.cap_nnz) return;
  if (range == 0 && lane == 0) prm.offset[sum->total[kRows]] = sum->total[kNnz];

  const auto start_of = [&](int c, uint32_t r) -> unsigned long long {
    return r < prm.num_ranges
               ? static_cast<unsigned long long>(prm.counts[c * prm.num_ranges + r]) +
                     prm.tile_base[c * prm.num_tiles + r / kScanTile]
               : sum->total[c];
  };
  const uint32_t b_rows = static_cast<uint32_t>(start_of(kRows, range));
  const uint32_t b_nnz = static_cast<uint32_t>(start_of(kNnz, range));
  const uint32_t n_rows = static_cast<uint32_t>(start_of(kRows, range + 1u)) - b_rows;
  const uint32_t n_value = static_cast<uint32_t>(start_of(kNnz, range + 1u)) - b_nnz;
  const Text txt = {prm.text, prm.size};
  const uint32_t lo = range * prm.range_bytes;
  const uint16_t *q16 = sp.queues + static_cast<size_t>(range) * sp.queue_stride;
  asm volatile("" : "+l"(q16));     /* kept in registers, see CollectRanges */
  const uint32_t o_row_at = sp.cap_pool, o_row_nnz = sp.cap_pool + sp.cap_rows, o_meta = sp.cap_pool + 2u * sp.cap_rows;
  const bool labelled = prm.label_column == 0;
  IdxT *const index = reinterpret_cast<IdxT *>(prm.index);

  /* values: queue entries three trips ahead, their text requested from L2 two trips ahead */
  const uint8_t *const text_lo = prm.text + lo;
  const uint32_t last_rel = prm.size - 1u - lo;       /* prefetches stay inside the chunk */
  const auto value_at = [&](uint32_t k) -> uint32_t { return k < n_value ? q16[k] : 0u; };
  uint32_t p_value = value_at(lane), a_value = value_at(lane + 32u), b_value = value_at(lane + 64u);
  for (uint32_t i = lane; i < n_value; i += 32u) {
    const uint32_t c_value = value_at(i + 96u);
    if (i + 64u < n_value) {
      asm volatile("prefetch.global.L1 [%0];" ::"l"(text_lo + b_value));
      asm volatile("prefetch.global.L1 [%0];" ::"l"(text_lo + min(b_value + 20u, last_rel)));
    }
    prm.value[b_nnz + i] = ToFloatFast(txt, lo + p_value);
    p_value = a_value;
    a_value = b_value;
    b_value = c_value;
  }
  for (uint32_t i = lane; i < n_rows; i += 32u) {
    prm.label[b_rows + i] = labelled ? ToFloatFast(txt, lo + q16[o_row_at + i]) : 0.0f;
    prm.offset[b_rows + i] = static_cast<unsigned long long>(b_nnz + q16[o_row_nnz + i]);
  }
  /* column ids: the j-th value of a row has id j (csv_parser.h:84-88).  The values in front of
     the range's first row start continue the row that was open when the range began: their ids go
     on from the values that row holds in the ranges in front (walk back to the nearest range
     with a row start). */
  const uint32_t head_len = n_rows != 0u ? q16[o_row_nnz] : n_value;
  if (head_len != 0u) {
    uint32_t open = 0;
    for (uint32_t r = range; r != 0u;) {
      --r;
      const uint16_t *m = sp.queues + static_cast<size_t>(r) * sp.queue_stride + o_meta;
      open += m[0];
      if (m[1] != 0u) break;
    }
    for (uint32_t j = lane; j < head_len; j += 32u) index[b_nnz + j] = static_cast<IdxT>(open + j);
  }
  if (n_value >= 16u * n_rows) {
    for (uint32_t r = 0; r < n_rows; ++r) {
      const uint32_t s = q16[o_row_nnz + r];
      const uint32_t e = r + 1u < n_rows ? q16[o_row_nnz + r + 1u] : n_value;
      for (uint32_t j = s + lane; j < e; j += 32u) index[b_nnz + j] = static_cast<IdxT>(j - s);
    }
  } else {
    /* short rows: every value finds its row by bisection over the row starts */
    for (uint32_t i = head_len + lane; i < n_value; i += 32u) {
      uint32_t a = 0, b = n_rows;     /* rows [0, a) start at or before value i */
      while (a < b) {
        const uint32_t m = (a + b) >> 1;
        if (q16[o_row_nnz + m] <= i) a = m + 1u; else b = m;
      }
      index[b_nnz + i] = static_cast<IdxT>(i - q16[o_row_nnz + a - 1u]);
    }
  }
}

}  // namespace b200p
#endif  // DMLC_CORE_B200_SPLIT_KERNELS_CUH_
