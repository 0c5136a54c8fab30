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
 * LibSVM and LibFM only (CSV keeps the one-launch kernel: its pass 1 was a popcount anyway).
 * A range whose lines span more than 64 KiB, or hold more tokens or rows than its queues, raises
 * kErrFallback and the chunk goes through the two-pass kernels.
 */
#ifndef DMLC_CORE_B200_SPLIT_KERNELS_CUH_
#define DMLC_CORE_B200_SPLIT_KERNELS_CUH_

#include "text_kernels.cuh"

namespace b200p {

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
constexpr int kSplitWarps = 8;
constexpr int kSplitThreads = kSplitWarps * 32;

/* resident CTAs per SM the register budget of CollectRanges is held to.  Measured on B200 (500 MB
   per shape, ms per launch, LibSVM / LibFM u64 / Criteo): 4 CTAs (64 registers) 0.586 / 0.666 /
   0.617, 5 CTAs (48) 0.642 / 0.710 / 0.664, 6 CTAs (40) 0.662 / 0.775 / 0.686 -- the spills cost
   more instructions than the extra warps hide. */
#ifndef B200P_COLLECT_CTAS
#define B200P_COLLECT_CTAS 4
#endif
template <int FMT, typename IdxT, bool QID>
__global__ void __launch_bounds__(kSplitThreads, B200P_COLLECT_CTAS) CollectRanges(const SplitParams sp) {
  __shared__ uint32_t rings[kSplitWarps][kTokenQueue];
  __shared__ __align__(16) uint8_t lut[256];
  const WalkParams &prm = sp.walk;
  for (int c = threadIdx.x; c < 256; c += kSplitThreads) lut[c] = static_cast<uint8_t>(ClassEntry<FMT>(c));
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const uint32_t range = blockIdx.x * kSplitWarps + warp;
  if (range >= prm.num_ranges) return;
  uint32_t *const tokens = rings[warp];
  const Text txt = {prm.text, prm.size};
  uint16_t *const q16 = sp.queues + static_cast<size_t>(range) * sp.queue_stride;
  const SplitOffsets at = SplitLayout<FMT>(sp.cap_pool, sp.cap_rows);

  const uint32_t lo = range * prm.range_bytes;
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
    const uint32_t room_pool = sp.cap_pool - 64u, room_rows = sp.cap_rows - 32u, room_field = sp.cap_pool / 2u - 32u;
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
      if (base + kStepBytes - lo > 0xffffu) {
        /* positions are stored as 16-bit offsets from `lo`: a line that carries the range past
           64 KiB is left to the two-pass kernels */
        give_up = true;
        break;
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
        tokens[slot] = first4 + b + SeparatorKind(__funnelshift_r(p_lo, p_mid, b) & 0xffu);
        ++slot;
      }
      bits = s1;
      while (bits != 0u) {
        const uint32_t b = __ffs(bits) - 1;
        bits &= bits - 1u;
        tokens[slot] = first4 + 32u + b + SeparatorKind(__funnelshift_r(p_mid, p_hi, b) & 0xffu);
        ++slot;
      }
      fill += __shfl_sync(0xffffffffu, incl, 31);
      __syncwarp();

      /* full groups of 32 tokens (all of them once the owned lines are finished) */
      uint32_t head = 0;
      const uint32_t groups = done ? (fill + 31u) >> 5 : fill >> 5;
      for (uint32_t g = 0; g < groups; ++g) {
        const uint32_t count = min(fill - head, 32u);
        if (n_index + n_value > room_pool || n_rows > room_rows || (FMT == kLibFM && n_index > room_field)) {
          give_up = true;
          dense = true;
          done = true;
          break;
        }
        const GroupRoles r = TokenRoles<FMT, QID>(txt, tokens, head, count, lane, carry_pos, carry_first);
        /* one 16-bit store per token, into the queue of its role; the queues' running ends are
           kept as offsets into the range's block */
        const uint16_t rel = static_cast<uint16_t>(r.at - lo);
        if ((r.m_index & me) != 0u) q16[n_index + __popc(r.m_index & lt)] = rel;
        if ((r.m_value & me) != 0u) q16[v_end - __popc(r.m_value & lt)] = rel;
        if (r.m_label != 0u) {           /* a row starts in this group */
          if ((r.m_label & me) != 0u) {
            const uint32_t k = n_rows + __popc(r.m_label & lt);
            q16[at.row_at + k] = rel;
            q16[at.row_nnz + k] = static_cast<uint16_t>(n_index + __popc(r.m_index & lt));
          }
          n_rows += __popc(r.m_label);
          if (r.m_weight != 0u) {
            if ((r.m_weight & me) != 0u) q16[at.weight + n_weight + __popc(r.m_weight & lt)] = rel;
            n_weight += __popc(r.m_weight);
          }
        } else if (r.m_weight != 0u) {   /* label and weight fell into different groups */
          if ((r.m_weight & me) != 0u) q16[at.weight + n_weight + __popc(r.m_weight & lt)] = rel;
          n_weight += __popc(r.m_weight);
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
          if ((r.m_index & me) != 0u) q16[at.field + n_index + __popc(r.m_index & lt)] = static_cast<uint16_t>(prev_at - lo);
          carry_at = __shfl_sync(0xffffffffu, r.at, static_cast<int>(count) - 1);
        }
        if (FMT == kLibSVM && r.m_qid != 0u) {
          if ((r.m_qid & me) != 0u) q16[at.qid + n_qid + __popc(r.m_qid & lt)] = static_cast<uint16_t>(r.qid_at - lo);
          n_qid += __popc(r.m_qid);
        }
        n_index += __popc(r.m_index);
        const uint32_t nv = __popc(r.m_value);
        n_value += nv;
        v_end -= nv;
        carry_pos = r.next_pos;
        carry_first = r.next_first;
        head += count;
      }
      if (head != 0u && !give_up) {
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
template <int FMT, typename IdxT>
__global__ void __launch_bounds__(kSplitThreads) ConvertRanges(const SplitParams sp) {
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
  const uint16_t *const q16 = sp.queues + static_cast<size_t>(range) * sp.queue_stride;
  const SplitOffsets at = SplitLayout<FMT>(sp.cap_pool, sp.cap_rows);
  IdxT *const index = reinterpret_cast<IdxT *>(prm.index);

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
  bool neg = false;
  const uint32_t sweep = max(n[kNnz], n[kValues]);
  const uint8_t *const text_lo = prm.text + lo;
  const auto prefetch = [&](uint32_t rel) {
    const uint8_t *p = text_lo + rel;
    asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
    asm volatile("prefetch.global.L1 [%0];" ::"l"(p + 20));
  };
  uint32_t p_index = 0, p_field = 0, p_value = 0;
  if (lane < n[kNnz]) {
    p_index = q16[lane];
    if constexpr (FMT == kLibFM) p_field = q16[at.field + lane];
  }
  if (lane < n[kValues]) p_value = q16[sp.cap_pool - 1u - lane];
  for (uint32_t i = lane; i < sweep; i += 32u) {
    uint32_t nx_index = 0, nx_field = 0, nx_value = 0;
    const uint32_t j = i + 32u;
    if (j < n[kNnz]) {
      nx_index = q16[j];
      if constexpr (FMT == kLibFM) nx_field = q16[at.field + j];
      prefetch(FMT == kLibFM ? nx_field : nx_index);
    }
    if (j < n[kValues]) {
      nx_value = q16[sp.cap_pool - 1u - j];
      prefetch(nx_value);
    }
    if (i < n[kNnz]) {
      if constexpr (FMT == kLibFM) {
        reinterpret_cast<IdxT *>(prm.field)[b[kNnz] + i] =
            static_cast<IdxT>(ToUintFast<sizeof(IdxT) == 8>(txt, lo + p_field, &neg));
      }
      index[b[kNnz] + i] = static_cast<IdxT>(ToUintFast<sizeof(IdxT) == 8>(txt, lo + p_index, &neg));
    }
    if (i < n[kValues]) prm.value[b[kValues] + i] = ToFloatFast(txt, lo + p_value);
    p_index = nx_index;
    p_field = nx_field;
    p_value = nx_value;
  }
  if (__any_sync(0xffffffffu, neg) && lane == 0) atomicOr(&prm.summary->error, static_cast<unsigned int>(kErrNegativeIndex));
}

}  // namespace b200p
#endif  // DMLC_CORE_B200_SPLIT_KERNELS_CUH_
