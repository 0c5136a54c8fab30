/*
 * Host side of the C ABI declared in include/dmlc_b200_parser.h: buffer management, kernel
 * launches and the copy/compute pipeline around the kernels of text_kernels.cuh.
 *
 * There is deliberately no CPU parsing code in this file or anywhere in the library: without a
 * usable GPU b200p_create fails with B200P_E_CUDA.
 */
#include <cuda_runtime.h>
#include <pthread.h>
#include <sys/mman.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <functional>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "../../include/dmlc_b200_parser.h"
#include "host_pool.h"
#include "text_kernels.cuh"
#include "tile_kernels.cuh"
#include "split_kernels.cuh"

namespace b200p {

__global__ void ZeroLabels(float *label, const ChunkSummary *summary, unsigned long long cap,
                           unsigned long long cap_nnz) {
  /* CSV rows without a label column keep label 0.0f (reference csv_parser.h:77).  Like pass 2,
     writes nothing when the arrays cannot hold the chunk. */
  const unsigned long long rows = summary->total[kRows];
  if (rows > cap || summary->total[kNnz] > cap_nnz) return;
  for (unsigned long long i = blockIdx.x * static_cast<unsigned long long>(blockDim.x) + threadIdx.x;
       i < rows; i += static_cast<unsigned long long>(gridDim.x) * blockDim.x) {
    label[i] = 0.0f;
  }
}

struct Arrays {
  unsigned long long *offset;
  float *label;
  float *weight;
  unsigned long long *qid;
  void *field;
  void *index;
  float *value;
  uint64_t cap_rows;  /* rows the row-sized arrays can hold (offset holds cap_rows + 1) */
  uint64_t cap_nnz;
  unsigned char *narrow; /* LibFM: kNarrowHeader bytes + one byte per field id (see NarrowFields) */
  void *block;        /* the one allocation all the arrays are carved from */
  Arrays() { std::memset(this, 0, sizeof(*this)); }
};

struct Slot {
  cudaStream_t stream;
  bool own_stream;
  uint8_t *h_text;   /* pinned staging (streaming path only) */
  uint8_t *d_text;
  uint64_t text_cap;
  uint32_t *d_counts;
  uint64_t counts_cap; /* in ranges */
  unsigned char *d_tiles; /* one-launch kernel: ticket + look-back status words */
  uint64_t tiles_cap;     /* in tiles */
  bool ran_tiles;         /* the chunk in flight went through ParseTiles */
  bool ran_split;         /* ... through CollectRanges / ConvertRanges */
  unsigned char *d_queues; /* 256-byte header (the give-up word) + the role queues of every range */
  uint64_t queues_cap;     /* bytes */
  uint32_t split_range, split_pool, split_rows, split_stride;   /* what the chunk in flight was collected with */
  ChunkSummary *d_summary;
  ChunkSummary *h_summary; /* pinned */
  Arrays dev;
  Arrays host; /* pinned mirrors (streaming path only) */
  const uint8_t *text; /* device pointer of the chunk in flight */
  uint64_t nbytes;
  bool busy;
  cudaEvent_t offsets_home; /* streaming path: offset[] has reached the host (created on first use) */
  unsigned char *h_narrow;  /* pinned landing place of the one-byte LibFM field ids */
  uint64_t narrow_cap;
  /* streaming path: what the completion thread has done with the chunk in flight */
  int stage;                /* 0 idle, 1 handed to the completion thread, 2 finished */
  int done_rc;              /* result of FinishParse (+ delivery) */
  bool delivered;           /* the arrays are already in the slot's pinned host mirrors */
  Slot() { std::memset(this, 0, sizeof(*this)); }
};

/* index[offset[r] + j] = j for the rows [r0, r1): what CSVParser::ParseBlock stores as the
   column id of the j-th non-label field of a row (reference csv_parser.h:84-88) */
template <typename T>
static void CsvColumnIds(const unsigned long long *offset, uint64_t r0, uint64_t r1, T *index,
                         const T *iota, uint64_t iota_len) {
  for (uint64_t r = r0; r < r1; ++r) {
    const uint64_t len = offset[r + 1] - offset[r];
    T *dst = index + offset[r];
    if (len <= iota_len) {
      std::memcpy(dst, iota, len * sizeof(T));
    } else {
      for (uint64_t j = 0; j < len; ++j) dst[j] = static_cast<T>(j);
    }
  }
}

}  // namespace b200p

using namespace b200p;

struct b200p_parser_s {
  b200p_config cfg;
  uint32_t range_bytes;
  int depth;
  std::vector<Slot> slots;   /* streaming path */
  Slot resident;             /* device-resident path */
  bool resident_pending;
  std::deque<int> in_flight; /* slot ids, oldest first */
  int lent_slot;             /* slot whose pinned input buffer was handed out, or -1 */
  int returned_slot;         /* slot whose host arrays the caller currently holds, or -1 */
  const Slot *last_block;    /* slot whose device arrays hold the block handed out last */
  double rows_per_byte, nnz_per_byte; /* densest chunk seen so far: sizes later allocations */
  std::atomic<uint64_t> launches;
  /* Streaming path, two threads: one may feed (b200p_input_buffer / b200p_submit) while another
     drains (b200p_next / b200p_pending / b200p_block_max).  `mu` guards the slot bookkeeping
     (slots' busy flags, in_flight, lent_slot, returned_slot, the density estimate); it is never
     held across a wait for a chunk in flight.  err_mu guards `error`, prof_mu one profiling bracket. */
  std::mutex mu, err_mu, prof_mu, deliver_mu;
  /* Streaming path, third thread (internal): finishes chunks in submission order as soon as
     their summary is home -- re-runs with larger arrays, hand-over to the two-pass kernels,
     error bits -- and, unless delivery is on demand or the results stay on the GPU, starts the
     device-to-host copies right away, so that they run while the consumer is still busy with
     the previous block (the reference's parse thread runs ahead of its consumer the same way,
     src/data/parser.h:73-82). */
  std::thread completer;
  bool completer_running, completer_stop;
  std::deque<int> to_complete;
  std::condition_variable cv_work, cv_done;
  int delivery;              /* B200P_DELIVER_* */
  unsigned long long *d_max; /* two words for b200p_block_max, allocated on first use */
  std::atomic<bool> saw_qids; /* a LibSVM chunk reported qids: later launches use the QID kernels */
  /* one-launch kernel (ParseTiles): used while the input suits it; a chunk it gives up on is
     parsed again by the two-pass kernels, and after kTileStrikes such chunks the handle stays
     with the two-pass kernels */
  std::atomic<bool> use_tiles, use_split;
  bool split_csv;
  std::atomic<int> last_path;   /* B200P_PATH_* of the chunk finished last */
  std::atomic<int> tile_fallbacks;
  uint32_t sub_bytes, over_bytes;
  std::atomic<uint32_t> pool_div;
  int host_threads;          /* b200p_set_host_threads; 0 = column ids always come from the GPU */
  HostPool *pool;            /* created on first use */
  std::vector<uint64_t> iota; /* 0, 1, 2, ... in the handle's index width */
  std::atomic<uint64_t> copied_h2d, copied_d2h;
  /* optional per-kernel timing (b200p_set_profiling): event pairs around each launch */
  bool profiling;
  struct Timed { int kind; cudaEvent_t start, stop; };
  std::vector<Timed> timed;
  double kernel_ms[3];
  uint64_t kernel_launches[3];
  std::string error;
};

static thread_local std::string g_create_error;

extern "C" {
static void CompleterLoop(b200p_handle h);
static void StopCompleter(b200p_handle h);
static int DeliverBlock(b200p_handle h, b200p::Slot *s, const b200p::Arrays &to, uint64_t base);
}

/* Every entry point works on its handle's device but leaves the caller's current device as it
   found it (the host program may be driving other GPUs from the same thread). */
class DeviceScope {
 public:
  explicit DeviceScope(int device) : previous_(-1), status_(cudaSuccess) {
    if (cudaGetDevice(&previous_) != cudaSuccess) previous_ = -1;
    if (previous_ != device) status_ = cudaSetDevice(device);
    else previous_ = -1;
  }
  ~DeviceScope() {
    if (previous_ >= 0) cudaSetDevice(previous_);
  }
  cudaError_t status() const { return status_; }
 private:
  int previous_;
  cudaError_t status_;
};

static void SetError(b200p_handle h, const char *msg) {
  std::lock_guard<std::mutex> g(h->err_mu);
  h->error = msg;
}

#define CUDA_TRY(h, expr)                                                              \
  do {                                                                                 \
    cudaError_t e_ = (expr);                                                           \
    if (e_ != cudaSuccess) {                                                           \
      char buf_[512];                                                                  \
      std::snprintf(buf_, sizeof(buf_), "CUDA error %s at %s:%d: %s", cudaGetErrorName(e_), \
                    __FILE__, __LINE__, cudaGetErrorString(e_));                       \
      SetError((h), buf_);                                                             \
      return B200P_E_CUDA;                                                             \
    }                                                                                  \
  } while (0)

/* profiling brackets: only active after b200p_set_profiling(h, 1) */
static void ProfBegin(b200p_handle h, int kind, cudaStream_t s) {
  if (!h->profiling) return;
  h->prof_mu.lock(); /* released by the matching ProfEnd: a bracket spans one launch */
  b200p_parser_s::Timed t;
  t.kind = kind;
  cudaEventCreate(&t.start);
  cudaEventCreate(&t.stop);
  cudaEventRecord(t.start, s);
  h->timed.push_back(t);
}
static void ProfEnd(b200p_handle h, cudaStream_t s) {
  if (!h->profiling) return;
  cudaEventRecord(h->timed.back().stop, s);
  h->prof_mu.unlock();
}
static void ProfCollect(b200p_handle h) {
  std::lock_guard<std::mutex> g(h->prof_mu);
  for (size_t i = 0; i < h->timed.size(); ++i) {
    float ms = 0.f;
    if (cudaEventSynchronize(h->timed[i].stop) == cudaSuccess &&
        cudaEventElapsedTime(&ms, h->timed[i].start, h->timed[i].stop) == cudaSuccess) {
      h->kernel_ms[h->timed[i].kind] += ms;
      h->kernel_launches[h->timed[i].kind] += 1;
    }
    cudaEventDestroy(h->timed[i].start);
    cudaEventDestroy(h->timed[i].stop);
  }
  h->timed.clear();
}

static int Fail(b200p_handle h, int code, const char *msg) {
  SetError(h, msg);
  return code;
}

/* ------------------------------------------------------------------ buffers ----------------- */
/*
 * Pinned host memory.  cudaHostAlloc pins 4 KiB page by 4 KiB page (measured on the B200 host:
 * 26 ms per 64 MiB = 2.4 GB/s, which made the first pass over a small file slower than the CPU
 * parser).  Buffers of 2 MiB and more are therefore carved from 2 MiB-aligned memory advised for
 * transparent huge pages and pinned with cudaHostRegister: 9 ms per 64 MiB with the first touch
 * included, same copy rates.  Small buffers, and hosts whose THP switch is "never", keep
 * cudaHostAlloc.  PinnedFree undoes either kind.
 */
namespace {
std::mutex g_registered_mu;
struct Registered {
  void *ptr;
  int device;        /* current when the block was pinned; made current again to unpin it */
  uint64_t bytes;    /* size of the block (a multiple of 2 MiB) */
};
std::vector<Registered> g_registered;   /* blocks obtained through the register path, in use */
/* Blocks a destroyed handle gave back, kept pinned for the next one: zero-filling, pinning and
   unpinning the 0.4-0.6 GB of a pipeline cost 35-50 ms each way (tools/first_pass.py), which is
   most of what a second parser of the same process spends on a file of a few hundred MB.
   B200P_PIN_CACHE_MB bounds what is kept (default 1024, 0 = nothing); blocks over 256 MiB (the
   arrays of an in-memory iterator) are never kept.  The blocks are pinned portably, so a
   handle on another device can take them. */
std::vector<Registered> g_pin_cache;
uint64_t g_pin_cache_bytes = 0;
uint64_t PinCacheLimit() {
  static const uint64_t limit = [] {
    const char *v = std::getenv("B200P_PIN_CACHE_MB");
    const long long mb = (v != NULL && v[0] != '\0') ? std::atoll(v) : 1024;
    return static_cast<uint64_t>(mb < 0 ? 0 : mb) << 20;
  }();
  return limit;
}

bool HugePagesOnOffer() {
  static const bool on = [] {
    const char *off = std::getenv("B200P_NO_HUGE_PAGES");
    if (off != NULL && off[0] != '\0' && off[0] != '0') return false;
    std::FILE *f = std::fopen("/sys/kernel/mm/transparent_hugepage/enabled", "r");
    if (f == NULL) return false;
    char line[128] = {0};
    const bool got = std::fgets(line, sizeof(line), f) != NULL;
    std::fclose(f);
    return got && std::strstr(line, "[never]") == NULL;
  }();
  return on;
}
}  // namespace

/* B200P_TRACE=1: where the set-up time of a handle went, printed by b200p_destroy */
static std::atomic<uint64_t> g_pinned_ns(0), g_pinned_bytes(0), g_device_ns(0), g_device_calls(0);
static uint64_t NowNs() {
  return static_cast<uint64_t>(std::chrono::duration_cast<std::chrono::nanoseconds>(
      std::chrono::steady_clock::now().time_since_epoch()).count());
}
struct AllocTimer {
  std::atomic<uint64_t> *into;
  uint64_t t0;
  explicit AllocTimer(std::atomic<uint64_t> *where) : into(where), t0(NowNs()) {}
  ~AllocTimer() { *into += NowNs() - t0; }
};

static cudaError_t DeviceAlloc(void **out, uint64_t bytes) {
  AllocTimer t(&g_device_ns);
  ++g_device_calls;
  return cudaMalloc(out, bytes);
}

static cudaError_t PinnedAlloc(void **out, uint64_t bytes) {
  AllocTimer t(&g_pinned_ns);
  g_pinned_bytes += bytes;
  const uint64_t kHuge = 2ull << 20;
  *out = NULL;
  if (bytes >= kHuge && HugePagesOnOffer()) {
    void *p = NULL;
    const uint64_t rounded = (bytes + kHuge - 1) / kHuge * kHuge;
    {
      /* the smallest kept block that fits and is not more than twice too large */
      std::lock_guard<std::mutex> g(g_registered_mu);
      size_t best = g_pin_cache.size();
      for (size_t i = 0; i < g_pin_cache.size(); ++i) {
        const uint64_t have = g_pin_cache[i].bytes;
        if (have >= rounded && have <= 2 * rounded &&
            (best == g_pin_cache.size() || have < g_pin_cache[best].bytes)) {
          best = i;
        }
      }
      if (best != g_pin_cache.size()) {
        const Registered r = g_pin_cache[best];
        g_pin_cache.erase(g_pin_cache.begin() + static_cast<long>(best));
        g_pin_cache_bytes -= r.bytes;
        g_registered.push_back(r);
        g_pinned_bytes -= bytes;      /* not a fresh allocation */
        *out = r.ptr;
        return cudaSuccess;
      }
    }
    if (posix_memalign(&p, kHuge, rounded) == 0) {
#ifdef MADV_HUGEPAGE
      madvise(p, rounded, MADV_HUGEPAGE);
#endif
      /* first touch is where the kernel zero-fills the pages (7 of the 9 ms per 64 MiB): spread
         it over a few threads before the pages are pinned */
      const unsigned hw = std::thread::hardware_concurrency();
      const uint64_t touchers = std::min<uint64_t>(std::min<uint64_t>(8, std::max(1u, hw / 2)), rounded / kHuge);
      if (touchers > 1) {
        std::vector<std::thread> crew;
        const uint64_t per = (rounded / touchers + kHuge - 1) / kHuge * kHuge;
        for (uint64_t i = 0; i < touchers; ++i) {
          crew.emplace_back([=] {
            volatile char *base = static_cast<volatile char *>(p);
            for (uint64_t at = per * i; at < std::min(rounded, per * (i + 1)); at += 4096) base[at] = 0;
          });
        }
        for (std::thread &th : crew) th.join();
      }
      if (cudaHostRegister(p, rounded, cudaHostRegisterPortable) == cudaSuccess) {
        Registered r = {p, 0, rounded};
        if (cudaGetDevice(&r.device) != cudaSuccess) r.device = 0;
        std::lock_guard<std::mutex> g(g_registered_mu);
        g_registered.push_back(r);
        *out = p;
        return cudaSuccess;
      }
      cudaGetLastError();   /* fall back to the allocator that always works */
      std::free(p);
    }
  }
  return cudaHostAlloc(out, std::max<uint64_t>(bytes, 1), cudaHostAllocDefault);
}

static void PinnedFree(void *p) {
  if (p == NULL) return;
  int device = -1;
  {
    std::lock_guard<std::mutex> g(g_registered_mu);
    for (size_t i = 0; i < g_registered.size(); ++i) {
      if (g_registered[i].ptr == p) {
        const Registered r = g_registered[i];
        g_registered.erase(g_registered.begin() + static_cast<long>(i));
        if (r.bytes <= (256ull << 20) && g_pin_cache_bytes + r.bytes <= PinCacheLimit()) {
          g_pin_cache.push_back(r);
          g_pin_cache_bytes += r.bytes;
          return;
        }
        device = r.device;
        break;
      }
    }
  }
  if (device < 0) {
    cudaFreeHost(p);
    return;
  }
  int current = -1;
  if (cudaGetDevice(&current) != cudaSuccess) current = -1;
  if (current != device) cudaSetDevice(device);
  cudaHostUnregister(p);
  if (current >= 0 && current != device) cudaSetDevice(current);
  std::free(p);
}

static void FreeDeviceArrays(Arrays *a) {
  cudaFree(a->block);
  *a = Arrays();
}
static void FreeHostArrays(Arrays *a) {
  PinnedFree(a->block);
  *a = Arrays();
}

/*
 * (Re)allocate so that `rows` rows and `nnz` entries fit; contents are not kept.  The seven
 * arrays of a set share one allocation (each starts on a 256-byte boundary): allocations are the
 * start-up cost of a handle, pinned ones by the byte and device ones by the call.  Caller has
 * synchronised the stream.
 */
static int EnsureArrays(b200p_handle h, Arrays *a, uint64_t rows, uint64_t nnz, bool pinned) {
  if (a->block != NULL && rows <= a->cap_rows && nnz <= a->cap_nnz) return B200P_OK;
  const int fmt = h->cfg.format;
  const uint64_t iw = h->cfg.index_bytes;
  const uint64_t cap_rows = rows > a->cap_rows || a->block == NULL ? rows + rows / 4 + 1024 : a->cap_rows;
  const uint64_t cap_nnz = nnz > a->cap_nnz || a->block == NULL ? nnz + nnz / 4 + 1024 : a->cap_nnz;
  const bool has_weight = fmt != B200P_FORMAT_CSV, has_qid = fmt == B200P_FORMAT_LIBSVM;
  const bool has_field = fmt == B200P_FORMAT_LIBFM;
  const auto pad = [](uint64_t bytes) { return (bytes + 255) / 256 * 256; };
  const uint64_t sizes[8] = {pad((cap_rows + 1) * 8), pad(cap_rows * 4), has_weight ? pad(cap_rows * 4) : 0,
                             has_qid ? pad(cap_rows * 8) : 0, has_field ? pad(cap_nnz * iw) : 0,
                             pad(cap_nnz * iw), pad(cap_nnz * 4),
                             has_field && !pinned ? pad(kNarrowHeader + cap_nnz) : 0};
  uint64_t total = 0;
  for (int i = 0; i < 8; ++i) total += sizes[i];
  if (pinned) PinnedFree(a->block); else cudaFree(a->block);
  *a = Arrays();
  void *block = NULL;
  CUDA_TRY(h, pinned ? PinnedAlloc(&block, total) : DeviceAlloc(&block, total));
  char *at = static_cast<char *>(block);
  void *parts[8];
  for (int i = 0; i < 8; ++i) {
    parts[i] = sizes[i] != 0 ? at : NULL;
    at += sizes[i];
  }
  a->block = block;
  a->offset = static_cast<unsigned long long *>(parts[0]);
  a->label = static_cast<float *>(parts[1]);
  a->weight = static_cast<float *>(parts[2]);
  a->qid = static_cast<unsigned long long *>(parts[3]);
  a->field = parts[4];
  a->index = parts[5];
  a->value = static_cast<float *>(parts[6]);
  a->narrow = static_cast<unsigned char *>(parts[7]);
  a->cap_rows = cap_rows;
  a->cap_nnz = cap_nnz;
  return B200P_OK;
}

/* Output capacity to provision for a chunk of `nbytes` before its counts are known. */
static void Provision(const b200p_handle h, uint64_t nbytes, uint64_t *rows, uint64_t *nnz) {
  if (h->nnz_per_byte > 0.0) {
    *rows = static_cast<uint64_t>(h->rows_per_byte * 1.25 * nbytes) + 1024;
    *nnz = static_cast<uint64_t>(h->nnz_per_byte * 1.25 * nbytes) + 1024;
  } else {
    *rows = nbytes / 32 + 1024; /* first chunk: a guess; pass 2 is re-run if it was too small */
    *nnz = nbytes / 4 + 1024;
  }
}

/* ------------------------------------------------------------------ launches ---------------- */
template <int FMT, typename IdxT>
static void LaunchWalks(const WalkParams &p, cudaStream_t s, bool emit, bool qids = false) {
  const unsigned grid = (p.num_ranges + kWarpsPerCta - 1) / kWarpsPerCta;
#ifdef B200P_CARVEOUT   /* tuning experiments only: shared-memory share of the L1/shared array, percent */
  static bool once = false;
  if (!once) {
    once = true;
    cudaFuncSetAttribute(WalkText<FMT, IdxT, true>, cudaFuncAttributePreferredSharedMemoryCarveout, B200P_CARVEOUT);
    cudaFuncSetAttribute(WalkText<FMT, IdxT, false>, cudaFuncAttributePreferredSharedMemoryCarveout, B200P_CARVEOUT);
  }
#endif
  if (FMT == kLibSVM && qids) {   /* the input has "qid:" on its rows: see TokenGroup */
    if (emit) {
      WalkText<FMT, IdxT, true, FMT == kLibSVM><<<grid, kCtaThreads, 0, s>>>(p);
    } else {
      WalkText<FMT, IdxT, false, FMT == kLibSVM><<<grid, kCtaThreads, 0, s>>>(p);
    }
  } else if (emit) {
    WalkText<FMT, IdxT, true><<<grid, kCtaThreads, 0, s>>>(p);
  } else {
    WalkText<FMT, IdxT, false><<<grid, kCtaThreads, 0, s>>>(p);
  }
}

static void LaunchWalk(const b200p_handle h, const WalkParams &p, cudaStream_t s, bool emit) {
  const bool wide = h->cfg.index_bytes == 8;
  switch (h->cfg.format) {
    case B200P_FORMAT_LIBSVM:
      if (wide) LaunchWalks<kLibSVM, uint64_t>(p, s, emit, h->saw_qids);
      else LaunchWalks<kLibSVM, uint32_t>(p, s, emit, h->saw_qids);
      break;
    case B200P_FORMAT_LIBFM:
      if (wide) LaunchWalks<kLibFM, uint64_t>(p, s, emit); else LaunchWalks<kLibFM, uint32_t>(p, s, emit);
      break;
    default:
      if (wide) LaunchWalks<kCSV, uint64_t>(p, s, emit); else LaunchWalks<kCSV, uint32_t>(p, s, emit);
      break;
  }
}

/* Bytes of text one warp owns. Large ranges amortise the two partial steps at the ends of a range
   (measured on the 1 GB CSV window: 8 KiB 1.20 ms, 16 KiB 1.14 ms, 32 KiB 1.13 ms per emit launch),
   small chunks need small ranges to fill 148 SMs x ~48 resident warps. */
static uint32_t RangeBytesFor(const b200p_handle h, uint64_t nbytes) {
  if (h->range_bytes != 0u) return h->range_bytes;
  if (nbytes >= (512ull << 20)) return 32768u;
  if (nbytes >= (128ull << 20)) return 16384u;
  if (nbytes >= (32ull << 20)) return 8192u;
  return 4096u;
}

static WalkParams MakeParams(const b200p_handle h, const Slot &s) {
  WalkParams p;
  std::memset(&p, 0, sizeof(p));
  p.text = s.text;
  p.size = static_cast<uint32_t>(s.nbytes);
  p.range_bytes = RangeBytesFor(h, s.nbytes);
  p.num_ranges = static_cast<uint32_t>((s.nbytes + p.range_bytes - 1) / p.range_bytes);
  p.label_column = h->cfg.label_column;
  p.four = 4u;
  p.sixteen = 16u;
  p.counts = s.d_counts;
  p.num_tiles = (p.num_ranges + kScanTile - 1) / kScanTile;
  p.tile_base = s.d_counts + static_cast<size_t>(kNumCounters) * p.num_ranges;
  p.summary = s.d_summary;
  p.offset = s.dev.offset;
  p.label = s.dev.label;
  p.weight = s.dev.weight;
  p.qid = s.dev.qid;
  p.field = s.dev.field;
  p.index = s.dev.index;
  p.value = s.dev.value;
  p.cap_rows = s.dev.cap_rows;
  p.cap_nnz = s.dev.cap_nnz;
  return p;
}


/* ---- the one-launch kernel (tile_kernels.cuh) ---- */
static const int kTileStrikes = 2;

static bool TilesSuit(const b200p_handle h) {
  /* CSV with the label in a later column needs a column id per field: two-pass kernels */
  return h->use_tiles && !(h->cfg.format == B200P_FORMAT_CSV && h->cfg.label_column > 0);
}

static TileParams MakeTileParams(const b200p_handle h, const Slot &s) {
  TileParams p;
  std::memset(&p, 0, sizeof(p));
  p.text = s.text;
  p.size = static_cast<uint32_t>(s.nbytes);
  p.sub_bytes = h->sub_bytes;
  p.over_bytes = h->cfg.format == B200P_FORMAT_CSV ? static_cast<uint32_t>(kStepBytes) : h->over_bytes;
  const uint64_t tile_bytes = static_cast<uint64_t>(kTileWarps) * p.sub_bytes;
  p.num_tiles = static_cast<uint32_t>((s.nbytes + tile_bytes - 1) / tile_bytes);
  p.label_column = h->cfg.label_column;
  p.four = 4u;
  p.sixteen = 16u;
  /* queue capacities per warp.  CSV: a sub-range cannot start more than sub_bytes / 2 fields.
     LibSVM / LibFM: the lines a warp owns may run into the overhang; text denser than about one
     token per four bytes sends the chunk to the two-pass kernels. */
  const uint32_t div = h->pool_div;      /* 4: ordinary text; 2: after a chunk proved denser (FinishParse) */
  p.cap_pool = h->cfg.format == B200P_FORMAT_CSV ? p.sub_bytes / 2u : std::max(256u, p.sub_bytes / div);
  p.cap_rows = std::max(64u, p.sub_bytes / (h->cfg.format == B200P_FORMAT_CSV ? 2u * div : 4u * div));
  p.scratch = s.d_tiles;
  p.summary = s.d_summary;
  p.offset = s.dev.offset;
  p.label = s.dev.label;
  p.weight = s.dev.weight;
  p.qid = s.dev.qid;
  p.field = s.dev.field;
  p.index = s.dev.index;
  p.value = s.dev.value;
  p.out_rows = s.dev.cap_rows;
  p.out_nnz = s.dev.cap_nnz;
  return p;
}

static uint64_t TileScratchBytes(uint64_t tiles) {
  return kTileHeaderBytes + tiles * 2ull * kStatusWords * sizeof(unsigned long long);
}

template <int FMT, typename IdxT, bool QID>
static cudaError_t LaunchTilesAs(const TileParams &p, cudaStream_t st) {
  const uint32_t smem = TileSmemBytes<FMT>(p);
  const cudaError_t e = cudaFuncSetAttribute(ParseTiles<FMT, IdxT, QID>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
  if (e != cudaSuccess) return e;
  ParseTiles<FMT, IdxT, QID><<<p.num_tiles, kTileThreads, smem, st>>>(p);
  return cudaGetLastError();
}

static cudaError_t LaunchTiles(int format, bool wide, bool qids, const TileParams &p, cudaStream_t st) {
  if (format == B200P_FORMAT_LIBSVM) {
    if (qids) return wide ? LaunchTilesAs<kLibSVM, uint64_t, true>(p, st) : LaunchTilesAs<kLibSVM, uint32_t, true>(p, st);
    return wide ? LaunchTilesAs<kLibSVM, uint64_t, false>(p, st) : LaunchTilesAs<kLibSVM, uint32_t, false>(p, st);
  }
  if (format == B200P_FORMAT_LIBFM) {
    return wide ? LaunchTilesAs<kLibFM, uint64_t, false>(p, st) : LaunchTilesAs<kLibFM, uint32_t, false>(p, st);
  }
  return wide ? LaunchTilesAs<kCSV, uint64_t, false>(p, st) : LaunchTilesAs<kCSV, uint32_t, false>(p, st);
}

/* clear the look-back scratch and the summary, one launch, summary to the host */
static int EnqueueTiles(b200p_handle h, Slot *s) {
  const TileParams p = MakeTileParams(h, *s);
  CUDA_TRY(h, cudaMemsetAsync(s->d_tiles, 0, TileScratchBytes(p.num_tiles), s->stream));
  CUDA_TRY(h, cudaMemsetAsync(s->d_summary, 0, sizeof(ChunkSummary), s->stream));
  ProfBegin(h, B200P_KERNEL_EMIT, s->stream);
  const cudaError_t e = LaunchTiles(h->cfg.format, h->cfg.index_bytes == 8, h->saw_qids, p, s->stream);
  ProfEnd(h, s->stream);
  ++h->launches;
  CUDA_TRY(h, e);
  CUDA_TRY(h, cudaMemcpyAsync(s->h_summary, s->d_summary, sizeof(ChunkSummary),
                              cudaMemcpyDeviceToHost, s->stream));
  return B200P_OK;
}

static int ProvisionArrays(b200p_handle h, Slot *s);

/* ---- collect / convert as two launches (split_kernels.cuh): LibSVM and LibFM ---- */
static bool SplitSuits(const b200p_handle h) {
  if (!h->use_split || !h->use_tiles) return false;
  /* CSV with the label in a later column needs a column id per field: two-pass kernels.
     B200P_SPLIT_CSV=0 keeps CSV on the one-launch kernel. */
  if (h->cfg.format == B200P_FORMAT_CSV) return h->split_csv && h->cfg.label_column <= 0;
  return true;
}

static SplitParams MakeSplitParams(const b200p_handle h, const Slot &s) {
  SplitParams sp;
  std::memset(&sp, 0, sizeof(sp));
  sp.walk = MakeParams(h, s);
  sp.walk.range_bytes = s.split_range;
  sp.walk.num_ranges = static_cast<uint32_t>((s.nbytes + s.split_range - 1) / s.split_range);
  sp.walk.num_tiles = (sp.walk.num_ranges + kScanTile - 1) / kScanTile;
  sp.walk.tile_base = s.d_counts + static_cast<size_t>(kNumCounters) * sp.walk.num_ranges;
  sp.queues = reinterpret_cast<uint16_t *>(s.d_queues + 256);
  sp.queue_stride = s.split_stride;
  sp.cap_pool = s.split_pool;
  sp.cap_rows = s.split_rows;
  sp.give_up = reinterpret_cast<uint32_t *>(s.d_queues);
  return sp;
}

template <int FMT, typename IdxT>
static void LaunchCollect(const SplitParams &sp, cudaStream_t st, bool qids) {
  const unsigned grid = (sp.walk.num_ranges + kSplitWarps - 1) / kSplitWarps;
  if (FMT == kLibSVM && qids) {
    CollectRanges<FMT, IdxT, FMT == kLibSVM><<<grid, kSplitThreads, 0, st>>>(sp);
  } else {
    CollectRanges<FMT, IdxT, false><<<grid, kSplitThreads, 0, st>>>(sp);
  }
}
template <int FMT, typename IdxT>
static void LaunchConvert(const SplitParams &sp, cudaStream_t st) {
  const unsigned grid = (sp.walk.num_ranges + kSplitWarps - 1) / kSplitWarps;
  ConvertRanges<FMT, IdxT><<<grid, kSplitThreads, 0, st>>>(sp);
}

static int EnqueueConvert(b200p_handle h, Slot *s) {
  const SplitParams sp = MakeSplitParams(h, *s);
  const bool wide = h->cfg.index_bytes == 8;
  ProfBegin(h, B200P_KERNEL_EMIT, s->stream);
  if (h->cfg.format == B200P_FORMAT_CSV) {
    const unsigned grid = (sp.walk.num_ranges + kSplitWarps - 1) / kSplitWarps;
    if (wide) ConvertFieldRanges<uint64_t><<<grid, kSplitThreads, 0, s->stream>>>(sp);
    else ConvertFieldRanges<uint32_t><<<grid, kSplitThreads, 0, s->stream>>>(sp);
  } else if (h->cfg.format == B200P_FORMAT_LIBSVM) {
    if (wide) LaunchConvert<kLibSVM, uint64_t>(sp, s->stream); else LaunchConvert<kLibSVM, uint32_t>(sp, s->stream);
  } else {
    if (wide) LaunchConvert<kLibFM, uint64_t>(sp, s->stream); else LaunchConvert<kLibFM, uint32_t>(sp, s->stream);
  }
  ProfEnd(h, s->stream);
  ++h->launches;
  CUDA_TRY(h, cudaGetLastError());
  CUDA_TRY(h, cudaMemcpyAsync(s->h_summary, s->d_summary, sizeof(ChunkSummary),
                              cudaMemcpyDeviceToHost, s->stream));
  return B200P_OK;
}

static int EnqueueSplit(b200p_handle h, Slot *s) {
  const bool libfm = h->cfg.format == B200P_FORMAT_LIBFM, csv = h->cfg.format == B200P_FORMAT_CSV;
  const uint32_t range_bytes = std::min<uint32_t>(RangeBytesFor(h, s->nbytes), 32768u);   /* 16-bit positions */
  const uint64_t ranges = (s->nbytes + range_bytes - 1) / range_bytes;
  const uint32_t div = h->pool_div;
  s->split_range = range_bytes;
  if (csv) {
    s->split_pool = range_bytes / 2u;                              /* a range cannot start more fields */
    s->split_rows = std::max(64u, range_bytes / (2u * div));       /* rows of 8 (4 once dense) bytes and more */
    s->split_stride = SplitCsvEntries(s->split_pool, s->split_rows);
  } else {
    s->split_pool = std::max(256u, range_bytes / div);
    s->split_rows = std::max(64u, range_bytes / (4u * div));
    s->split_stride = libfm ? SplitQueueEntries<kLibFM>(s->split_pool, s->split_rows)
                            : SplitQueueEntries<kLibSVM>(s->split_pool, s->split_rows);
  }
  if (ranges > s->counts_cap) {
    CUDA_TRY(h, cudaStreamSynchronize(s->stream));
    cudaFree(s->d_counts);
    s->d_counts = NULL;
    s->counts_cap = 0;
    const uint64_t cap = ranges + ranges / 4 + 64;
    CUDA_TRY(h, DeviceAlloc(reinterpret_cast<void **>(&s->d_counts),
                           (cap + cap / kScanTile + 2) * kNumCounters * sizeof(uint32_t)));
    s->counts_cap = cap;
  }
  const uint64_t need = 256 + ranges * s->split_stride * sizeof(uint16_t);
  if (need > s->queues_cap) {
    CUDA_TRY(h, cudaStreamSynchronize(s->stream));
    cudaFree(s->d_queues);
    s->d_queues = NULL;
    s->queues_cap = 0;
    const uint64_t cap = need + need / 8;
    CUDA_TRY(h, DeviceAlloc(reinterpret_cast<void **>(&s->d_queues), cap));
    s->queues_cap = cap;
  }
  const int rc = ProvisionArrays(h, s);
  if (rc != B200P_OK) return rc;
  const SplitParams sp = MakeSplitParams(h, *s);
  const bool wide = h->cfg.index_bytes == 8;
  CUDA_TRY(h, cudaMemsetAsync(s->d_queues, 0, 256, s->stream));
  ProfBegin(h, B200P_KERNEL_COUNT, s->stream);
  if (csv) {
    const unsigned grid = (sp.walk.num_ranges + kSplitWarps - 1) / kSplitWarps;
    if (wide) CollectFieldRanges<uint64_t><<<grid, kSplitThreads, 0, s->stream>>>(sp);
    else CollectFieldRanges<uint32_t><<<grid, kSplitThreads, 0, s->stream>>>(sp);
  } else if (libfm) {
    if (wide) LaunchCollect<kLibFM, uint64_t>(sp, s->stream, false); else LaunchCollect<kLibFM, uint32_t>(sp, s->stream, false);
  } else {
    if (wide) LaunchCollect<kLibSVM, uint64_t>(sp, s->stream, h->saw_qids); else LaunchCollect<kLibSVM, uint32_t>(sp, s->stream, h->saw_qids);
  }
  ProfEnd(h, s->stream);
  ProfBegin(h, B200P_KERNEL_SCAN, s->stream);
  ScanTiles<<<sp.walk.num_tiles, kScanTile, 0, s->stream>>>(sp.walk.counts, sp.walk.num_ranges, sp.walk.tile_base, sp.walk.num_tiles);
  ScanFinish<<<1, kScanTile, 0, s->stream>>>(sp.walk.tile_base, sp.walk.num_tiles, sp.walk.num_ranges, s->d_summary, sp.give_up);
  ProfEnd(h, s->stream);
  h->launches += 3;
  CUDA_TRY(h, cudaGetLastError());
  s->ran_tiles = false;
  s->ran_split = true;
  return EnqueueConvert(h, s);
}

/* pass 2 (+ the CSV label clear), then the summary travels to the host */
static int EnqueueEmit(b200p_handle h, Slot *s) {
  const WalkParams p = MakeParams(h, *s);
  if (h->cfg.format == B200P_FORMAT_CSV) {
    ZeroLabels<<<296, 256, 0, s->stream>>>(s->dev.label, s->d_summary, s->dev.cap_rows, s->dev.cap_nnz);
    ++h->launches;
  }
  ProfBegin(h, B200P_KERNEL_EMIT, s->stream);
  LaunchWalk(h, p, s->stream, true);
  ProfEnd(h, s->stream);
  ++h->launches;
  CUDA_TRY(h, cudaGetLastError());
  CUDA_TRY(h, cudaMemcpyAsync(s->h_summary, s->d_summary, sizeof(ChunkSummary),
                              cudaMemcpyDeviceToHost, s->stream));
  return B200P_OK;
}

/* output arrays for a chunk whose counts are not known yet */
static int ProvisionArrays(b200p_handle h, Slot *s) {
  uint64_t want_rows, want_nnz;
  Provision(h, s->nbytes, &want_rows, &want_nnz);
  if (want_rows > s->dev.cap_rows || want_nnz > s->dev.cap_nnz || s->dev.offset == NULL) {
    CUDA_TRY(h, cudaStreamSynchronize(s->stream));
    const int rc = EnsureArrays(h, &s->dev, want_rows, want_nnz, false);
    if (rc != B200P_OK) return rc;
  }
  return B200P_OK;
}

/* pass 1 + scan + (speculative) pass 2 for the chunk described by s->text / s->nbytes */
static int EnqueueTwoPass(b200p_handle h, Slot *s) {
  s->ran_tiles = false;
  s->ran_split = false;
  const uint32_t range_bytes = RangeBytesFor(h, s->nbytes);
  const uint64_t ranges = (s->nbytes + range_bytes - 1) / range_bytes;
  if (ranges > s->counts_cap) {
    CUDA_TRY(h, cudaStreamSynchronize(s->stream));
    cudaFree(s->d_counts);
    s->d_counts = NULL;
    s->counts_cap = 0;
    const uint64_t cap = ranges + ranges / 4 + 64;
    /* per-range counters followed by the per-tile bases of the scan */
    CUDA_TRY(h, DeviceAlloc(reinterpret_cast<void **>(&s->d_counts),
                           (cap + cap / kScanTile + 2) * kNumCounters * sizeof(uint32_t)));
    s->counts_cap = cap;
  }
  const int rc = ProvisionArrays(h, s);
  if (rc != B200P_OK) return rc;
  const WalkParams p = MakeParams(h, *s);
  ProfBegin(h, B200P_KERNEL_COUNT, s->stream);
  LaunchWalk(h, p, s->stream, false);
  ProfEnd(h, s->stream);
  ProfBegin(h, B200P_KERNEL_SCAN, s->stream);
  ScanTiles<<<p.num_tiles, kScanTile, 0, s->stream>>>(p.counts, p.num_ranges, p.tile_base, p.num_tiles);
  ScanFinish<<<1, kScanTile, 0, s->stream>>>(p.tile_base, p.num_tiles, p.num_ranges, s->d_summary, NULL);
  ProfEnd(h, s->stream);
  h->launches += 3;
  CUDA_TRY(h, cudaGetLastError());
  return EnqueueEmit(h, s);
}

/* the kernels for the chunk described by s->text / s->nbytes: one launch where the input suits
   ParseTiles, the two-pass walk otherwise */
static int EnqueueParse(b200p_handle h, Slot *s) {
  if (s->nbytes == 0) {
    std::memset(s->h_summary, 0, sizeof(ChunkSummary));
    return B200P_OK;
  }
  if (SplitSuits(h)) return EnqueueSplit(h, s);
  if (!TilesSuit(h)) return EnqueueTwoPass(h, s);
  const uint64_t tile_bytes = static_cast<uint64_t>(kTileWarps) * h->sub_bytes;
  const uint64_t tiles = (s->nbytes + tile_bytes - 1) / tile_bytes;
  if (tiles > s->tiles_cap) {
    CUDA_TRY(h, cudaStreamSynchronize(s->stream));
    cudaFree(s->d_tiles);
  cudaFree(s->d_queues);
    s->d_tiles = NULL;
    s->tiles_cap = 0;
    const uint64_t cap = tiles + tiles / 4 + 16;
    CUDA_TRY(h, DeviceAlloc(reinterpret_cast<void **>(&s->d_tiles), TileScratchBytes(cap)));
    s->tiles_cap = cap;
  }
  const int rc = ProvisionArrays(h, s);
  if (rc != B200P_OK) return rc;
  s->ran_tiles = true;
  s->ran_split = false;
  return EnqueueTiles(h, s);
}

/* Wait for the chunk, re-run pass 2 with larger arrays if the provision was too small, and
   translate device error bits. On success the device arrays of `s` hold the chunk. */
static int FinishParse(b200p_handle h, Slot *s) {
  if (s->nbytes == 0) return B200P_OK;
  CUDA_TRY(h, cudaStreamSynchronize(s->stream));
  const ChunkSummary &sum = *s->h_summary;
  if ((s->ran_tiles || s->ran_split) && (sum.error & kErrFallback) != 0u && (sum.error & kErrInternal) == 0u) {
    /* the one-launch kernel met something it does not take (a line longer than its overhang,
       text denser than its queues): this chunk goes through the two-pass kernels, and a handle
       that keeps meeting such chunks stops trying */
    if ((sum.error & kErrDense) != 0u && h->pool_div > 2u) {
      h->pool_div = 2u;      /* short tokens or short rows: larger queues (fewer resident CTAs) from the next chunk on */
    } else if (++h->tile_fallbacks >= kTileStrikes) {
      h->use_tiles = false;
    }
    const int rc = EnqueueTwoPass(h, s);
    if (rc != B200P_OK) return rc;
    CUDA_TRY(h, cudaStreamSynchronize(s->stream));
  }
  if (sum.total[kRows] > s->dev.cap_rows || sum.total[kNnz] > s->dev.cap_nnz) {
    const int rc = EnsureArrays(h, &s->dev, sum.total[kRows], sum.total[kNnz], false);
    if (rc != B200P_OK) return rc;
    const int rc2 = s->ran_tiles ? EnqueueTiles(h, s) : (s->ran_split ? EnqueueConvert(h, s) : EnqueueEmit(h, s));
    if (rc2 != B200P_OK) return rc2;
    CUDA_TRY(h, cudaStreamSynchronize(s->stream));
  }
  h->last_path = s->ran_tiles ? B200P_PATH_ONE_LAUNCH : (s->ran_split ? B200P_PATH_COLLECT_CONVERT : B200P_PATH_TWO_PASS);
  if (sum.total[kQids] != 0) h->saw_qids = true;
  {
    std::lock_guard<std::mutex> g(h->mu);
    h->rows_per_byte = std::max(h->rows_per_byte, static_cast<double>(sum.total[kRows]) / s->nbytes);
    h->nnz_per_byte = std::max(h->nnz_per_byte, static_cast<double>(sum.total[kNnz]) / s->nbytes);
  }
  if (sum.error & kErrInternal) {
    return Fail(h, B200P_E_CUDA, "internal consistency check failed (checked build): a queue or output index left its bounds");
  }
  if (sum.error & kErrNegativeIndex) {
    return Fail(h, B200P_E_PARSE, "Check failed: sign == true (negative value in an unsigned index/field)");
  }
  if (sum.error & kErrValueCount) {
    return Fail(h, B200P_E_PARSE,
                "Check failed: offset.back() == value.size() || value.size() == 0");
  }
  return B200P_OK;
}

static void FillBlock(const b200p_handle h, const Slot &s, const Arrays &a, bool on_device,
                      b200p_rowblock *out) {
  std::memset(out, 0, sizeof(*out));
  out->index_bytes = h->cfg.index_bytes;
  out->on_device = on_device ? 1 : 0;
  out->text_bytes = s.nbytes;
  if (s.nbytes == 0) return;
  const ChunkSummary &sum = *s.h_summary;
  out->rows = sum.total[kRows];
  out->nnz = sum.total[kNnz];
  out->n_weight = sum.total[kWeights];
  out->n_qid = sum.total[kQids];
  out->n_value = sum.total[kValues];
  out->n_field = h->cfg.format == B200P_FORMAT_LIBFM ? out->nnz : 0;
  if (out->rows == 0) return;
  out->offset = reinterpret_cast<const uint64_t *>(a.offset);
  out->label = a.label;
  out->weight = out->n_weight ? a.weight : NULL;
  out->qid = out->n_qid ? reinterpret_cast<const uint64_t *>(a.qid) : NULL;
  out->field = out->n_field ? a.field : NULL;
  out->index = out->nnz ? a.index : NULL;
  out->value = out->n_value ? a.value : NULL;
}

static void DestroySlot(Slot *s) {
  if (s->stream != NULL) cudaStreamSynchronize(s->stream);
  PinnedFree(s->h_text);
  PinnedFree(s->h_narrow);
  cudaFree(s->d_text);
  cudaFree(s->d_counts);
  cudaFree(s->d_tiles);
  cudaFree(s->d_queues);
  cudaFree(s->d_summary);
  cudaFreeHost(s->h_summary);
  FreeDeviceArrays(&s->dev);
  FreeHostArrays(&s->host);
  if (s->offsets_home != NULL) cudaEventDestroy(s->offsets_home);
  if (s->own_stream && s->stream != NULL) cudaStreamDestroy(s->stream);
  *s = Slot();
}

static int InitSlot(b200p_handle h, Slot *s, cudaStream_t stream) {
  if (stream != NULL) {
    s->stream = stream;
    s->own_stream = false;
  } else {
    CUDA_TRY(h, cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
    s->own_stream = true;
  }
  CUDA_TRY(h, cudaMalloc(reinterpret_cast<void **>(&s->d_summary), sizeof(ChunkSummary)));
  CUDA_TRY(h, cudaHostAlloc(reinterpret_cast<void **>(&s->h_summary), sizeof(ChunkSummary), cudaHostAllocDefault));
  std::memset(s->h_summary, 0, sizeof(ChunkSummary));
  return B200P_OK;
}

/* host.index of a CSV chunk from host.offset (already on the host), spread over the pool */
static void RebuildCsvIndex(b200p_handle h, const Arrays &to, uint64_t rows) {
  const bool wide = h->cfg.index_bytes == 8;
  const uint64_t kIota = 4096;
  if (h->iota.empty()) {
    h->iota.resize(wide ? kIota : kIota / 2);
    if (wide) {
      for (uint64_t j = 0; j < kIota; ++j) h->iota[j] = j;
    } else {
      uint32_t *narrow = reinterpret_cast<uint32_t *>(h->iota.data());
      for (uint64_t j = 0; j < kIota; ++j) narrow[j] = static_cast<uint32_t>(j);
    }
  }
  if (h->pool == NULL) h->pool = new HostPool(h->host_threads);
  const int parts = static_cast<int>(std::min<uint64_t>(rows, 4ull * h->pool->threads()));
  const unsigned long long *offset = to.offset;
  void *index = to.index;
  const uint64_t *iota = h->iota.data();
  h->pool->Run(parts, [=](int part) {
    const uint64_t r0 = rows * part / parts, r1 = rows * (part + 1) / parts;
    if (wide) {
      CsvColumnIds(offset, r0, r1, static_cast<uint64_t *>(index), iota, kIota);
    } else {
      CsvColumnIds(offset, r0, r1, static_cast<uint32_t *>(index),
                   reinterpret_cast<const uint32_t *>(iota), kIota);
    }
  });
}

/* host.field of a LibFM chunk from the one-byte ids in host.narrow, spread over the pool */
static void WidenFields(b200p_handle h, const Slot *s, const Arrays &dest, uint64_t nnz) {
  if (h->pool == NULL) h->pool = new HostPool(h->host_threads);
  const int parts = static_cast<int>(std::min<uint64_t>((nnz + 65535) / 65536, 4ull * h->pool->threads()));
  const unsigned char *from = s->h_narrow + kNarrowHeader;
  void *to = dest.field;
  const bool wide = h->cfg.index_bytes == 8;
  h->pool->Run(parts, [=](int part) {
    const uint64_t lo = nnz * part / parts, hi = nnz * (part + 1) / parts;
    if (wide) {
      uint64_t *dst = static_cast<uint64_t *>(to);
      for (uint64_t i = lo; i < hi; ++i) dst[i] = from[i];
    } else {
      uint32_t *dst = static_cast<uint32_t *>(to);
      for (uint64_t i = lo; i < hi; ++i) dst[i] = from[i];
    }
  });
}

/* ------------------------------------------------------------------ C ABI ------------------- */
extern "C" {

#ifdef B200P_CHECKED
const char *b200p_version(void) { return "dmlc_b200 0.1 sm_100a checked"; }
#else
const char *b200p_version(void) { return "dmlc_b200 0.1 sm_100a"; }
#endif

int b200p_current_device(void) {
  int d = -1;
  if (cudaGetDevice(&d) != cudaSuccess) {
    cudaGetLastError();
    return -1;
  }
  return d;
}

int b200p_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

const char *b200p_last_error(b200p_handle h) {
  if (h == NULL) return g_create_error.c_str();
  static thread_local std::string mine; /* a copy: the other thread may fail while this one reads */
  std::lock_guard<std::mutex> g(h->err_mu);
  mine = h->error;
  return mine.c_str();
}

int b200p_create(const b200p_config *config, b200p_handle *out) {
  if (out == NULL) return B200P_E_INVALID;
  *out = NULL;
  if (config == NULL || config->format < 0 || config->format > 2 ||
      (config->index_bytes != 4 && config->index_bytes != 8)) {
    g_create_error = "b200p_create: format must be 0..2 and index_bytes 4 or 8";
    return B200P_E_INVALID;
  }
  b200p_handle h = new (std::nothrow) b200p_parser_s();
  if (h == NULL) return B200P_E_INVALID;
  h->cfg = *config;
  if (h->cfg.chunk_bytes == 0) h->cfg.chunk_bytes = 64ull << 20;
  if (h->cfg.chunk_bytes > B200P_MAX_CHUNK_BYTES) h->cfg.chunk_bytes = B200P_MAX_CHUNK_BYTES;
  /* 0 = choose per chunk (RangeBytesFor) */
  uint32_t rb = config->range_bytes > 0 ? static_cast<uint32_t>(config->range_bytes) : 0u;
  if (rb != 0u) rb = std::max<uint32_t>(kStepBytes, (rb + kStepBytes - 1) / kStepBytes * kStepBytes);
  h->range_bytes = rb;
  h->depth = config->depth > 0 ? std::min(config->depth, 8) : 3;
  h->slots.reserve(static_cast<size_t>(h->depth) + 2);
  h->lent_slot = -1;
  h->returned_slot = -1;
  h->completer_running = false;
  h->completer_stop = false;
  h->delivery = B200P_DELIVER_EAGER;
  if (h->cfg.out_residency != B200P_OUT_DEVICE) h->cfg.out_residency = B200P_OUT_HOST;
  h->last_block = NULL;
  h->resident_pending = false;
  h->rows_per_byte = h->nnz_per_byte = 0.0;
  h->launches = 0;
  h->d_max = NULL;
  h->saw_qids = false;
  {
    /* B200P_TILES=0 keeps a process on the two-pass kernels (comparisons, tests of that path);
       B200P_SUB_BYTES / B200P_OVER_BYTES are tuning knobs of the one-launch kernel */
    const char *off = std::getenv("B200P_TILES");
    h->use_tiles = !(off != NULL && off[0] == '0');
    const char *off2 = std::getenv("B200P_SPLIT");    /* B200P_SPLIT=0: LibSVM / LibFM through ParseTiles as well */
    h->use_split = !(off2 != NULL && off2[0] == '0');
    const char *off3 = std::getenv("B200P_SPLIT_CSV");
    h->split_csv = !(off3 != NULL && off3[0] == '0');
    h->tile_fallbacks = 0;
    h->last_path = -1;
    uint32_t sub = 2048u, over = 4096u;
    const char *e1 = std::getenv("B200P_SUB_BYTES"), *e2 = std::getenv("B200P_OVER_BYTES");
    if (e1 != NULL && std::atoi(e1) > 0) sub = static_cast<uint32_t>(std::atoi(e1));
    if (e2 != NULL && std::atoi(e2) > 0) over = static_cast<uint32_t>(std::atoi(e2));
    if (rb != 0u) sub = rb;     /* config.range_bytes: bytes of text one warp owns, in both kernels */
    sub = std::min<uint32_t>(4096u, std::max<uint32_t>(kStepBytes, sub / kStepBytes * kStepBytes));
    over = std::min<uint32_t>(16384u, std::max<uint32_t>(kStepBytes, (over + kStepBytes - 1) / kStepBytes * kStepBytes));
    h->sub_bytes = sub;
    h->over_bytes = over;
    const char *e3 = std::getenv("B200P_POOL_DIV");
    h->pool_div = (e3 != NULL && std::atoi(e3) >= 2 && std::atoi(e3) <= 16) ? static_cast<uint32_t>(std::atoi(e3)) : 4u;
  }
  h->pool = NULL;
  h->copied_h2d = h->copied_d2h = 0;
  {
    /* default: a quarter of the host's cores, at most four, shared fairly between the ranks of
       one host when a launcher says how many there are (B200P_HOST_THREADS overrides) */
    const unsigned hw = std::thread::hardware_concurrency();
    const char *local = std::getenv("LOCAL_WORLD_SIZE");
    const unsigned ranks = (local != NULL && std::atoi(local) > 0) ? static_cast<unsigned>(std::atoi(local)) : 1u;
    int n = static_cast<int>(std::min(4u, std::max(1u, hw / 4u / ranks)));
    const char *env = std::getenv("B200P_HOST_THREADS");
    if (env != NULL && env[0] != '\0') n = std::atoi(env);
    h->host_threads = std::max(0, std::min(n, 64));
  }
  h->profiling = false;
  for (int k = 0; k < 3; ++k) { h->kernel_ms[k] = 0.0; h->kernel_launches[k] = 0; }

  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count <= 0 || config->device < 0 || config->device >= count) {
    char buf[256];
    std::snprintf(buf, sizeof(buf),
                  "b200p_create: no usable CUDA device %d (%s); this library has no CPU path",
                  config->device, e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0 or ordinal out of range");
    g_create_error = buf;
    delete h;
    return B200P_E_CUDA;
  }
  DeviceScope on_device(config->device);
  e = on_device.status();
  int rc = B200P_OK;
  if (e != cudaSuccess) {
    g_create_error = std::string("b200p_create: cudaSetDevice failed: ") + cudaGetErrorString(e);
    delete h;
    return B200P_E_CUDA;
  }
  rc = InitSlot(h, &h->resident, static_cast<cudaStream_t>(config->stream));
  if (rc != B200P_OK) {
    g_create_error = h->error;
    DestroySlot(&h->resident);
    delete h;
    return rc;
  }
  *out = h;
  return B200P_OK;
}

void b200p_destroy(b200p_handle h) {
  if (h == NULL) return;
  StopCompleter(h);
  DeviceScope on_device(h->cfg.device);
  ProfCollect(h);
  for (size_t i = 0; i < h->slots.size(); ++i) DestroySlot(&h->slots[i]);
  DestroySlot(&h->resident);
  cudaFree(h->d_max);
  delete h->pool;
  delete h;
  if (std::getenv("B200P_TRACE") != NULL) {
    std::fprintf(stderr, "b200p trace: pinned allocations %.1f ms for %.0f MB, device allocations %.1f ms in %llu calls (process totals)\n",
                 g_pinned_ns.load() * 1e-6, g_pinned_bytes.load() * 1e-6, g_device_ns.load() * 1e-6,
                 static_cast<unsigned long long>(g_device_calls.load()));
  }
}

uint64_t b200p_launch_count(b200p_handle h) { return h != NULL ? h->launches.load() : 0; }

int b200p_last_path(b200p_handle h) { return h != NULL ? h->last_path.load() : -1; }

int b200p_set_profiling(b200p_handle h, int on) {
  if (h == NULL) return B200P_E_INVALID;
  DeviceScope on_device(h->cfg.device);
  ProfCollect(h);
  h->profiling = on != 0;
  for (int k = 0; k < 3; ++k) { h->kernel_ms[k] = 0.0; h->kernel_launches[k] = 0; }
  return B200P_OK;
}

int b200p_kernel_time(b200p_handle h, int kernel, double *total_ms, uint64_t *launches) {
  if (h == NULL || kernel < 0 || kernel > 2) return B200P_E_INVALID;
  DeviceScope on_device(h->cfg.device);
  ProfCollect(h); /* waits for the recorded events */
  if (total_ms != NULL) *total_ms = h->kernel_ms[kernel];
  if (launches != NULL) *launches = h->kernel_launches[kernel];
  return B200P_OK;
}

/* ---- device-resident path ---- */
int b200p_parse_device_async(b200p_handle h, const void *d_text, uint64_t nbytes) {
  if (h == NULL) return B200P_E_INVALID;
  if (nbytes > B200P_MAX_CHUNK_BYTES) return Fail(h, B200P_E_INVALID, "chunk larger than B200P_MAX_CHUNK_BYTES");
  if ((reinterpret_cast<uintptr_t>(d_text) & 15u) != 0) return Fail(h, B200P_E_INVALID, "device chunk must be 16-byte aligned");
  DeviceScope on_device(h->cfg.device);
  CUDA_TRY(h, on_device.status());
  if (h->resident_pending) {
    /* one chunk per finish: the handle has one set of result arrays for this path, a second
       chunk would overwrite the first one's rows and its error bits would never be looked at */
    return Fail(h, B200P_E_AGAIN, "b200p_parse_device_async: a chunk is pending, call b200p_parse_device_finish first");
  }
  Slot *s = &h->resident;
  s->text = static_cast<const uint8_t *>(d_text);
  s->nbytes = nbytes;
  h->resident_pending = true;
  const int rc = EnqueueParse(h, s);
  if (rc != B200P_OK) h->resident_pending = false;
  return rc;
}

int b200p_parse_device_finish(b200p_handle h, b200p_rowblock *out) {
  if (h == NULL || out == NULL) return B200P_E_INVALID;
  if (!h->resident_pending) return Fail(h, B200P_E_EMPTY, "b200p_parse_device_finish without a pending chunk");
  DeviceScope on_device(h->cfg.device);
  CUDA_TRY(h, on_device.status());
  h->resident_pending = false;
  const int rc = FinishParse(h, &h->resident);
  if (rc != B200P_OK) return rc;
  FillBlock(h, h->resident, h->resident.dev, true, out);
  h->last_block = &h->resident;
  return B200P_OK;
}

int b200p_parse_device(b200p_handle h, const void *d_text, uint64_t nbytes, b200p_rowblock *out) {
  const int rc = b200p_parse_device_async(h, d_text, nbytes);
  if (rc != B200P_OK) return rc;
  return b200p_parse_device_finish(h, out);
}

int b200p_block_max(b200p_handle h, uint64_t *max_index, uint64_t *max_field) {
  if (h == NULL) return B200P_E_INVALID;
  if (max_index != NULL) *max_index = 0;
  if (max_field != NULL) *max_field = 0;
  if (h->last_block == NULL) return Fail(h, B200P_E_EMPTY, "b200p_block_max: no block has been returned yet");
  const Slot *s = h->last_block;
  if (s->nbytes == 0) return B200P_OK;
  const uint64_t nnz = s->h_summary->total[kNnz];
  if (nnz == 0) return B200P_OK;
  DeviceScope on_device(h->cfg.device);
  CUDA_TRY(h, on_device.status());
  /* two result words of its own, allocated once */
  if (h->d_max == NULL) CUDA_TRY(h, cudaMalloc(reinterpret_cast<void **>(&h->d_max), 2 * sizeof(unsigned long long)));
  CUDA_TRY(h, cudaMemsetAsync(h->d_max, 0, 2 * sizeof(unsigned long long), s->stream));
  const unsigned grid = static_cast<unsigned>(std::min<uint64_t>((nnz + 255) / 256, 148 * 8));
  const bool wide = h->cfg.index_bytes == 8;
  if (wide) {
    MaxOfArray<uint64_t><<<grid, 256, 0, s->stream>>>(static_cast<const uint64_t *>(s->dev.index), nnz, h->d_max);
    if (h->cfg.format == B200P_FORMAT_LIBFM)
      MaxOfArray<uint64_t><<<grid, 256, 0, s->stream>>>(static_cast<const uint64_t *>(s->dev.field), nnz, h->d_max + 1);
  } else {
    MaxOfArray<uint32_t><<<grid, 256, 0, s->stream>>>(static_cast<const uint32_t *>(s->dev.index), nnz, h->d_max);
    if (h->cfg.format == B200P_FORMAT_LIBFM)
      MaxOfArray<uint32_t><<<grid, 256, 0, s->stream>>>(static_cast<const uint32_t *>(s->dev.field), nnz, h->d_max + 1);
  }
  h->launches += h->cfg.format == B200P_FORMAT_LIBFM ? 2 : 1;
  CUDA_TRY(h, cudaGetLastError());
  unsigned long long host[2] = {0ull, 0ull};
  CUDA_TRY(h, cudaMemcpyAsync(host, h->d_max, sizeof(host), cudaMemcpyDeviceToHost, s->stream));
  CUDA_TRY(h, cudaStreamSynchronize(s->stream));
  if (max_index != NULL) *max_index = host[0];
  if (max_field != NULL) *max_field = host[1];
  return B200P_OK;
}

int b200p_fetch(b200p_handle h, void *host_dst, const void *device_src, uint64_t nbytes) {
  if (h == NULL || (nbytes != 0 && (host_dst == NULL || device_src == NULL))) return B200P_E_INVALID;
  if (nbytes == 0) return B200P_OK;
  DeviceScope on_device(h->cfg.device);
  CUDA_TRY(h, on_device.status());
  CUDA_TRY(h, cudaMemcpyAsync(host_dst, device_src, nbytes, cudaMemcpyDeviceToHost, h->resident.stream));
  CUDA_TRY(h, cudaStreamSynchronize(h->resident.stream));
  h->copied_d2h += nbytes;
  return B200P_OK;
}

int b200p_set_host_threads(b200p_handle h, int threads) {
  if (h == NULL || threads < 0 || threads > 64) return B200P_E_INVALID;
  if (h->pool != NULL && h->pool->threads() != threads) {
    delete h->pool;
    h->pool = NULL;
  }
  h->host_threads = threads;
  return B200P_OK;
}

int b200p_copied_bytes(b200p_handle h, uint64_t *host_to_device, uint64_t *device_to_host) {
  if (h == NULL) return B200P_E_INVALID;
  if (host_to_device != NULL) *host_to_device = h->copied_h2d.load();
  if (device_to_host != NULL) *device_to_host = h->copied_d2h.load();
  return B200P_OK;
}

/* ---- streaming path ---- */
static int FreeSlotIndex(b200p_handle h) {
  /* slots are created on demand up to depth; the one whose arrays the caller still holds
     (returned_slot) is not reusable until the next b200p_next */
  for (size_t i = 0; i < h->slots.size(); ++i) {
    if (!h->slots[i].busy && static_cast<int>(i) != h->returned_slot) return static_cast<int>(i);
  }
  /* depth in flight + one held by the caller + one being filled by its reader */
  if (static_cast<int>(h->slots.size()) < h->depth + 2) {
    h->slots.push_back(Slot());
    return static_cast<int>(h->slots.size()) - 1;
  }
  return -1;
}

static int PrepareSlot(b200p_handle h, int id, uint64_t nbytes) {
  Slot *s = &h->slots[id];
  if (s->stream == NULL) {
    const int rc = InitSlot(h, s, NULL);
    if (rc != B200P_OK) return rc;
  }
  const uint64_t want = std::max<uint64_t>(nbytes, h->cfg.chunk_bytes);
  if (want > s->text_cap) {
    CUDA_TRY(h, cudaStreamSynchronize(s->stream));
    PinnedFree(s->h_text);
    cudaFree(s->d_text);
    s->h_text = NULL;
    s->d_text = NULL;
    s->text_cap = 0;
    const uint64_t cap = (want + 4095) / 4096 * 4096;
    CUDA_TRY(h, PinnedAlloc(reinterpret_cast<void **>(&s->h_text), cap));
    CUDA_TRY(h, DeviceAlloc(reinterpret_cast<void **>(&s->d_text), cap));
    s->text_cap = cap;
  }
  return B200P_OK;
}

int b200p_input_buffer(b200p_handle h, void **host_ptr, uint64_t *capacity) {
  if (h == NULL || host_ptr == NULL) return B200P_E_INVALID;
  DeviceScope on_device(h->cfg.device);
  CUDA_TRY(h, on_device.status());
  std::lock_guard<std::mutex> g(h->mu);
  if (h->lent_slot < 0) {
    const int id = FreeSlotIndex(h);
    if (id < 0) return Fail(h, B200P_E_AGAIN, "all pipeline slots are in flight: call b200p_next");
    const int rc = PrepareSlot(h, id, 0);
    if (rc != B200P_OK) return rc;
    h->lent_slot = id;
  }
  *host_ptr = h->slots[h->lent_slot].h_text;
  if (capacity != NULL) *capacity = h->slots[h->lent_slot].text_cap;
  return B200P_OK;
}

int b200p_submit(b200p_handle h, const void *text, uint64_t nbytes) {
  if (h == NULL || (text == NULL && nbytes != 0)) return B200P_E_INVALID;
  if (nbytes > B200P_MAX_CHUNK_BYTES) return Fail(h, B200P_E_INVALID, "chunk larger than B200P_MAX_CHUNK_BYTES");
  DeviceScope on_device(h->cfg.device);
  CUDA_TRY(h, on_device.status());
  std::lock_guard<std::mutex> g(h->mu);
  int id = h->lent_slot;
  const bool in_place = id >= 0 && text == h->slots[id].h_text;
  if (id < 0) {
    id = FreeSlotIndex(h);
    if (id < 0) return Fail(h, B200P_E_AGAIN, "all pipeline slots are in flight: call b200p_next");
  }
  if (in_place && nbytes > h->slots[id].text_cap) return Fail(h, B200P_E_INVALID, "chunk exceeds the staging buffer");
  const void *src = text;
  if (!in_place) {
    const int rc = PrepareSlot(h, id, nbytes);
    if (rc != B200P_OK) return rc;
    /* memory the caller already pinned (cudaHostAlloc / cudaHostRegister) is copied to the
       device straight from where it lies; pageable memory goes through the staging buffer */
    bool caller_pinned = false;
    if (nbytes != 0) {
      cudaPointerAttributes attr;
      if (cudaPointerGetAttributes(&attr, text) == cudaSuccess) {
        caller_pinned = attr.type == cudaMemoryTypeHost;
      } else {
        cudaGetLastError();
      }
    }
    if (!caller_pinned && nbytes != 0) {
      std::memcpy(h->slots[id].h_text, text, nbytes);
      src = h->slots[id].h_text;
    }
  }
  h->lent_slot = -1;
  Slot *s = &h->slots[id];
  s->nbytes = nbytes;
  s->text = s->d_text;
  if (nbytes != 0) {
    CUDA_TRY(h, cudaMemcpyAsync(s->d_text, src, nbytes, cudaMemcpyHostToDevice, s->stream));
    h->copied_h2d += nbytes;
  }
  const int rc = EnqueueParse(h, s);
  if (rc != B200P_OK) return rc;
  s->busy = true;
  s->stage = 1;
  s->delivered = false;
  h->in_flight.push_back(id);
  h->to_complete.push_back(id);
  if (!h->completer_running) {
    h->completer_stop = false;
    h->completer = std::thread(CompleterLoop, h);
    h->completer_running = true;
  }
  h->cv_work.notify_one();
  return B200P_OK;
}

/* the completion thread: see b200p_parser_s */
static void CompleterLoop(b200p_handle h) {
  DeviceScope on_device(h->cfg.device);
  std::unique_lock<std::mutex> lk(h->mu);
  for (;;) {
    h->cv_work.wait(lk, [h] { return h->completer_stop || !h->to_complete.empty(); });
    if (h->completer_stop) break;
    const int id = h->to_complete.front();
    h->to_complete.pop_front();
    Slot *s = &h->slots[id];   /* slots never moves: reserved for depth + 2 entries at creation */
    const bool deliver = h->delivery == B200P_DELIVER_EAGER && h->cfg.out_residency == B200P_OUT_HOST;
    lk.unlock();
    int rc = FinishParse(h, s);
    bool delivered = false;
    if (rc == B200P_OK && deliver && s->nbytes != 0) {
      const ChunkSummary &sum = *s->h_summary;
      rc = EnsureArrays(h, &s->host, sum.total[kRows], sum.total[kNnz], true);
      if (rc == B200P_OK) rc = DeliverBlock(h, s, s->host, 0);
      delivered = rc == B200P_OK;
    }
    lk.lock();
    s->done_rc = rc;
    s->delivered = delivered;
    s->stage = 2;
    h->cv_done.notify_all();
  }
}

static void StopCompleter(b200p_handle h) {
  {
    std::lock_guard<std::mutex> g(h->mu);
    if (!h->completer_running) return;
    h->completer_stop = true;
  }
  h->cv_work.notify_all();
  h->completer.join();
  h->completer_running = false;
}

int b200p_set_delivery(b200p_handle h, int mode) {
  if (h == NULL || (mode != B200P_DELIVER_EAGER && mode != B200P_DELIVER_ON_DEMAND)) return B200P_E_INVALID;
  std::lock_guard<std::mutex> g(h->mu);
  h->delivery = mode;
  return B200P_OK;
}

int b200p_pending(b200p_handle h) {
  if (h == NULL) return 0;
  std::lock_guard<std::mutex> g(h->mu);
  return static_cast<int>(h->in_flight.size());
}

/*
 * The finished chunk of slot `s` into the host arrays `to` (the slot's own pinned mirrors for
 * b200p_next, the caller's memory for b200p_next_to): copies, the host-side rebuild of CSV
 * column ids / LibFM field ids, and a final synchronisation.  `base` is added to every offset.
 */
static int DeliverBlock(b200p_handle h, Slot *s, const Arrays &to, uint64_t base) {
  std::lock_guard<std::mutex> one_at_a_time(h->deliver_mu);   /* the host pool and its tables are shared */
  const ChunkSummary &sum = *s->h_summary;
  const uint64_t rows = sum.total[kRows], nnz = sum.total[kNnz];
  {
    const size_t iw = h->cfg.index_bytes;
    cudaStream_t st = s->stream;
    if (rows != 0) {
      /* Wide CSV rows: index[] is fully determined by offset[] (the j-th non-label field of a row
         has column id j, csv_parser.h:84-88), so those nnz * index_bytes bytes are not brought
         back over PCIe -- about half of this path's device-to-host traffic -- but written by a
         few host threads while the value[] copy is still on the link.  Rows shorter than 16
         entries on average are cheaper to copy than to rebuild row by row. */
      const bool rebuild = h->cfg.format == B200P_FORMAT_CSV && h->host_threads > 0 &&
                           nnz >= 16 * rows;
      /* LibFM field ids cross the link as one byte each and are widened to IndexType by the same
         host threads (NarrowFields); an id above 255 anywhere in the chunk brings the wide array
         over after all. */
      const bool narrow = h->cfg.format == B200P_FORMAT_LIBFM && h->host_threads > 0 && nnz >= 4096;
      uint64_t bytes = (rows + 1) * 8 + rows * 4;
      if (rebuild || narrow) {
        if (s->offsets_home == NULL) {
          CUDA_TRY(h, cudaEventCreateWithFlags(&s->offsets_home, cudaEventDisableTiming));
        }
      }
      if (narrow) {
        if (kNarrowHeader + nnz > s->narrow_cap) {
          PinnedFree(s->h_narrow);
          s->h_narrow = NULL;
          s->narrow_cap = 0;
          const uint64_t cap = kNarrowHeader + nnz + nnz / 4;
          CUDA_TRY(h, PinnedAlloc(reinterpret_cast<void **>(&s->h_narrow), cap));
          s->narrow_cap = cap;
        }
        CUDA_TRY(h, cudaMemsetAsync(s->dev.narrow, 0, 4, st));
        const unsigned grid = static_cast<unsigned>(std::min<uint64_t>((nnz + 255) / 256, 148 * 8));
        if (iw == 8) {
          NarrowFields<uint64_t><<<grid, 256, 0, st>>>(static_cast<const uint64_t *>(s->dev.field), nnz, s->dev.narrow);
        } else {
          NarrowFields<uint32_t><<<grid, 256, 0, st>>>(static_cast<const uint32_t *>(s->dev.field), nnz, s->dev.narrow);
        }
        ++h->launches;
        CUDA_TRY(h, cudaGetLastError());
        CUDA_TRY(h, cudaMemcpyAsync(s->h_narrow, s->dev.narrow, kNarrowHeader + nnz, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(h, cudaEventRecord(s->offsets_home, st));
        bytes += kNarrowHeader + nnz;
      }
      CUDA_TRY(h, cudaMemcpyAsync(to.offset, s->dev.offset, (rows + 1) * 8, cudaMemcpyDeviceToHost, st));
      if (rebuild) CUDA_TRY(h, cudaEventRecord(s->offsets_home, st));
      CUDA_TRY(h, cudaMemcpyAsync(to.label, s->dev.label, rows * 4, cudaMemcpyDeviceToHost, st));
      if (sum.total[kWeights] != 0) {
        CUDA_TRY(h, cudaMemcpyAsync(to.weight, s->dev.weight, sum.total[kWeights] * 4, cudaMemcpyDeviceToHost, st));
        bytes += sum.total[kWeights] * 4;
      }
      if (sum.total[kQids] != 0) {
        CUDA_TRY(h, cudaMemcpyAsync(to.qid, s->dev.qid, sum.total[kQids] * 8, cudaMemcpyDeviceToHost, st));
        bytes += sum.total[kQids] * 8;
      }
      if (nnz != 0) {
        if (!rebuild) {
          CUDA_TRY(h, cudaMemcpyAsync(to.index, s->dev.index, nnz * iw, cudaMemcpyDeviceToHost, st));
          bytes += nnz * iw;
        }
        if (h->cfg.format == B200P_FORMAT_LIBFM && !narrow) {
          CUDA_TRY(h, cudaMemcpyAsync(to.field, s->dev.field, nnz * iw, cudaMemcpyDeviceToHost, st));
          bytes += nnz * iw;
        }
      }
      if (sum.total[kValues] != 0) {
        CUDA_TRY(h, cudaMemcpyAsync(to.value, s->dev.value, sum.total[kValues] * 4, cudaMemcpyDeviceToHost, st));
        bytes += sum.total[kValues] * 4;
      }
      if (rebuild) {
        CUDA_TRY(h, cudaEventSynchronize(s->offsets_home));
        RebuildCsvIndex(h, to, rows);
      }
      if (narrow) {
        CUDA_TRY(h, cudaEventSynchronize(s->offsets_home));
        unsigned int overflow = 0;
        std::memcpy(&overflow, s->h_narrow, sizeof(overflow));
        if (overflow != 0u) {
          CUDA_TRY(h, cudaMemcpyAsync(to.field, s->dev.field, nnz * iw, cudaMemcpyDeviceToHost, st));
          bytes += nnz * iw;
        } else {
          WidenFields(h, s, to, nnz);
        }
      }
      CUDA_TRY(h, cudaStreamSynchronize(st));
      h->copied_d2h += bytes;
    }
    if (base != 0 && rows != 0) {
      for (uint64_t i = 0; i <= rows; ++i) to.offset[i] += base;
    }
  }
  return B200P_OK;
}

/* copy a block the completion thread already brought home into the caller's arrays */
static void CopyDelivered(const b200p_handle h, const Slot &s, const Arrays &to, uint64_t base) {
  const ChunkSummary &sum = *s.h_summary;
  const uint64_t rows = sum.total[kRows], nnz = sum.total[kNnz];
  const size_t iw = h->cfg.index_bytes;
  if (rows == 0) return;
  for (uint64_t i = 0; i <= rows; ++i) to.offset[i] = s.host.offset[i] + base;
  std::memcpy(to.label, s.host.label, rows * 4);
  if (sum.total[kWeights] != 0) std::memcpy(to.weight, s.host.weight, sum.total[kWeights] * 4);
  if (sum.total[kQids] != 0) std::memcpy(to.qid, s.host.qid, sum.total[kQids] * 8);
  if (nnz != 0) {
    std::memcpy(to.index, s.host.index, nnz * iw);
    if (h->cfg.format == B200P_FORMAT_LIBFM) std::memcpy(to.field, s.host.field, nnz * iw);
  }
  if (sum.total[kValues] != 0) std::memcpy(to.value, s.host.value, sum.total[kValues] * 4);
}

int b200p_next(b200p_handle h, b200p_rowblock *out) {
  if (h == NULL || out == NULL) return B200P_E_INVALID;
  DeviceScope on_device(h->cfg.device);
  CUDA_TRY(h, on_device.status());
  Slot *s = NULL;
  int rc = B200P_OK;
  bool delivered = false;
  {
    std::unique_lock<std::mutex> lk(h->mu);
    if (h->in_flight.empty()) return Fail(h, B200P_E_EMPTY, "b200p_next: nothing submitted");
    const int id = h->in_flight.front();
    h->in_flight.pop_front();
    s = &h->slots[id];   /* slots never moves: reserved for depth + 2 entries at creation */
    s->busy = false;
    /* the block handed out before this call is released and this one is held, in one step, so
       that a concurrent b200p_submit cannot pick the slot being finished here */
    h->returned_slot = id;
    h->last_block = NULL;
    h->cv_done.wait(lk, [s] { return s->stage == 2; });
    rc = s->done_rc;
    delivered = s->delivered;
    s->stage = 0;
  }
  if (rc != B200P_OK) return rc;
  if (h->cfg.out_residency == B200P_OUT_DEVICE) {
    /* the consumer lives on the GPU: the arrays stay where the kernels wrote them */
    FillBlock(h, *s, s->dev, true, out);
    h->last_block = s;
    return B200P_OK;
  }
  if (s->nbytes != 0 && !delivered) {
    const ChunkSummary &sum = *s->h_summary;
    const uint64_t rows = sum.total[kRows], nnz = sum.total[kNnz];
    rc = EnsureArrays(h, &s->host, rows, nnz, true);
    if (rc != B200P_OK) return rc;
    rc = DeliverBlock(h, s, s->host, 0);
    if (rc != B200P_OK) return rc;
  }
  FillBlock(h, *s, s->host, false, out);
  h->last_block = s;
  return B200P_OK;
}

int b200p_next_to(b200p_handle h, const b200p_destination *dst, b200p_rowblock *out) {
  if (h == NULL || dst == NULL || out == NULL) return B200P_E_INVALID;
  DeviceScope on_device(h->cfg.device);
  CUDA_TRY(h, on_device.status());
  if (h->cfg.out_residency != B200P_OUT_HOST) {
    return Fail(h, B200P_E_INVALID, "b200p_next_to: this handle keeps its results on the GPU (out_residency)");
  }
  Slot *s = NULL;
  int id = -1;
  int rc = B200P_OK;
  bool delivered = false;
  {
    std::unique_lock<std::mutex> lk(h->mu);
    if (h->in_flight.empty()) return Fail(h, B200P_E_EMPTY, "b200p_next_to: nothing submitted");
    id = h->in_flight.front();   /* stays queued until it is known to fit: only this thread pops */
    s = &h->slots[id];
    h->cv_done.wait(lk, [s] { return s->stage == 2; });
    rc = s->done_rc;
    delivered = s->delivered;
  }
  Arrays to;
  to.offset = reinterpret_cast<unsigned long long *>(dst->offset);
  to.label = dst->label;
  to.weight = dst->weight;
  to.qid = reinterpret_cast<unsigned long long *>(dst->qid);
  to.field = dst->field;
  to.index = dst->index;
  to.value = dst->value;
  std::memset(out, 0, sizeof(*out));
  if (rc == B200P_OK && s->nbytes != 0) {
    /* refusals leave the block queued and say what it holds */
    const ChunkSummary &sum = *s->h_summary;
    out->rows = sum.total[kRows];
    out->nnz = sum.total[kNnz];
    out->n_weight = sum.total[kWeights];
    out->n_qid = sum.total[kQids];
    out->n_value = sum.total[kValues];
    out->n_field = h->cfg.format == B200P_FORMAT_LIBFM ? out->nnz : 0;
    out->text_bytes = s->nbytes;
    if (sum.total[kRows] > dst->cap_rows || sum.total[kNnz] > dst->cap_nnz) {
      return Fail(h, B200P_E_CAPACITY, "b200p_next_to: the destination arrays are too small for the next block; "
                                       "it stays queued (the counts in *out say how much room it needs)");
    }
    if ((out->rows != 0 && (to.offset == NULL || to.label == NULL)) || (out->nnz != 0 && to.index == NULL) ||
        (out->n_value != 0 && to.value == NULL) || (out->n_weight != 0 && to.weight == NULL) ||
        (out->n_qid != 0 && to.qid == NULL) || (out->n_field != 0 && to.field == NULL)) {
      return Fail(h, B200P_E_INVALID, "b200p_next_to: a destination array the next block needs is NULL; "
                                      "it stays queued (the counts in *out say which)");
    }
  }
  {
    std::lock_guard<std::mutex> g(h->mu);
    h->in_flight.pop_front();
    s->busy = false;
    /* the slot stays held, as after b200p_next: its device arrays feed the copies below and
       b200p_block_max until the next block is taken */
    h->returned_slot = id;
    h->last_block = NULL;
  }
  s->stage = 0;
  if (rc != B200P_OK) return rc;
  if (s->nbytes != 0) {
    if (delivered) {
      CopyDelivered(h, *s, to, dst->offset_base);   /* brought home before the caller said where to */
    } else {
      rc = DeliverBlock(h, s, to, dst->offset_base);
      if (rc != B200P_OK) return rc;
    }
  }
  FillBlock(h, *s, to, false, out);
  h->last_block = s;           /* b200p_block_max still works: the device arrays are intact */
  return B200P_OK;
}

int b200p_host_alloc(b200p_handle h, uint64_t bytes, void **ptr) {
  if (h == NULL || ptr == NULL) return B200P_E_INVALID;
  DeviceScope on_device(h->cfg.device);
  CUDA_TRY(h, on_device.status());
  CUDA_TRY(h, PinnedAlloc(ptr, bytes));
  return B200P_OK;
}

void b200p_host_free(void *ptr) { PinnedFree(ptr); }


/* ------------------------------------------------------------------------------------------
 * b200p_parse_device_noalloc: the same count -> scan -> emit parse as b200p_parse_device, for
 * hosts that must not allocate GPU memory while they run.
 *
 * Who needs this: a parser embedded in a process whose HBM is already owned by somebody else's
 * arena -- an XGBoost/RMM pool or a deep-learning framework's caching allocator that reserved
 * the whole card at start-up -- or one that runs next to live NCCL collectives, where a
 * cudaMalloc / cudaFree can stall every stream of the device or simply fail because nothing is
 * left outside the pool.  Such a caller carves its text buffer and its CSR arrays out of its own
 * pool and hands them in; the only other memory the kernels need is their scratch (per-range
 * counters and the chunk summary), and that lives in the fixed array below, which the CUDA
 * module reserves once per GPU when the library is loaded and keeps until the process exits.
 *
 * Limit: kNoAllocMaxRanges ranges of kNoAllocRangeBytes, i.e. chunks up to 256 MiB; the array is
 * about 640 KiB per GPU.  Larger chunks are refused, not truncated.
 *
 * Shared state: every call on a GPU uses the same array and finds the previous call's counters
 * in it.  Pass 1 rewrites every counter it will read and the summary is cleared at the start of
 * each call, so calls issued one after another are independent -- but two calls whose kernels
 * overlap on one GPU (different streams, different threads) race on the array and corrupt each
 * other.  Serialise them.
 *
 * First call on a device: resolves the device address of the array with cudaGetSymbolAddress
 * (the host-side name `g_noalloc_scratch` is only a stand-in; handing it to a kernel would make
 * the kernel touch unrelated memory or fault).  Make that first call outside any stream capture.
 * ------------------------------------------------------------------------------------------ */
namespace {
constexpr uint32_t kNoAllocMaxRanges = 32768u;
constexpr uint32_t kNoAllocRangeBytes = 8192u;
constexpr uint64_t kNoAllocMaxBytes = static_cast<uint64_t>(kNoAllocMaxRanges) * kNoAllocRangeBytes;
constexpr int kNoAllocMaxDevices = 64;

struct NoAllocScratch {
  uint32_t counts[kNumCounters * kNoAllocMaxRanges];
  uint32_t tile_base[kNumCounters * (kNoAllocMaxRanges / kScanTile)];
  ChunkSummary summary;
};

/* file-scope scratch: one instance per GPU that loads this module, resident until exit */
static __device__ __align__(256) unsigned char g_noalloc_scratch[sizeof(NoAllocScratch)];

static NoAllocScratch *g_noalloc_address[kNoAllocMaxDevices]; /* device addresses, per ordinal */
static pthread_mutex_t g_noalloc_mutex = PTHREAD_MUTEX_INITIALIZER;

static void NoAllocSay(char *errbuf, uint64_t errlen, const char *msg) {
  if (errbuf != NULL && errlen != 0) std::snprintf(errbuf, errlen, "%s", msg);
}

/* device address of the scratch on `device` (which must be current); set up once per device */
static cudaError_t NoAllocScratchFor(int device, NoAllocScratch **out) {
  cudaError_t e = cudaSuccess;
  pthread_mutex_lock(&g_noalloc_mutex);
  if (g_noalloc_address[device] == NULL) {
    void *ptr = NULL;
    e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
    if (e == cudaSuccess) g_noalloc_address[device] = static_cast<NoAllocScratch *>(ptr);
  }
  *out = g_noalloc_address[device];
  pthread_mutex_unlock(&g_noalloc_mutex);
  return e;
}

static void LaunchWalkPlain(int format, bool wide, const WalkParams &p, cudaStream_t s, bool emit) {
  if (format == B200P_FORMAT_LIBSVM) {
    if (wide) LaunchWalks<kLibSVM, uint64_t>(p, s, emit); else LaunchWalks<kLibSVM, uint32_t>(p, s, emit);
  } else if (format == B200P_FORMAT_LIBFM) {
    if (wide) LaunchWalks<kLibFM, uint64_t>(p, s, emit); else LaunchWalks<kLibFM, uint32_t>(p, s, emit);
  } else {
    if (wide) LaunchWalks<kCSV, uint64_t>(p, s, emit); else LaunchWalks<kCSV, uint32_t>(p, s, emit);
  }
}
}  // namespace

int b200p_parse_device_noalloc(const b200p_noalloc_request *rq, b200p_rowblock *out,
                               char *errbuf, uint64_t errlen) {
  if (errbuf != NULL && errlen != 0) errbuf[0] = '\0';
  if (rq == NULL || out == NULL) return B200P_E_INVALID;
  std::memset(out, 0, sizeof(*out));
  if (rq->format < 0 || rq->format > 2 || (rq->index_bytes != 4 && rq->index_bytes != 8) ||
      rq->device < 0 || rq->device >= kNoAllocMaxDevices) {
    NoAllocSay(errbuf, errlen, "b200p_parse_device_noalloc: bad format / index_bytes / device");
    return B200P_E_INVALID;
  }
  if (rq->nbytes > kNoAllocMaxBytes) {
    if (errbuf != NULL && errlen != 0) {
      std::snprintf(errbuf, errlen,
                    "b200p_parse_device_noalloc: chunk of %llu bytes exceeds the fixed scratch, which "
                    "covers at most %llu bytes per call; split the chunk after a line end or use "
                    "b200p_parse_device",
                    static_cast<unsigned long long>(rq->nbytes),
                    static_cast<unsigned long long>(kNoAllocMaxBytes));
    }
    return B200P_E_INVALID;
  }
  if ((reinterpret_cast<uintptr_t>(rq->d_text) & 15u) != 0) {
    NoAllocSay(errbuf, errlen, "b200p_parse_device_noalloc: d_text must be 16-byte aligned");
    return B200P_E_INVALID;
  }
  const bool need_weight = rq->format != B200P_FORMAT_CSV;
  const bool need_qid = rq->format == B200P_FORMAT_LIBSVM;
  const bool need_field = rq->format == B200P_FORMAT_LIBFM;
  if (rq->nbytes != 0 &&
      (rq->d_text == NULL || rq->d_offset == NULL || rq->d_label == NULL || rq->d_index == NULL ||
       rq->d_value == NULL || (need_weight && rq->d_weight == NULL) ||
       (need_qid && rq->d_qid == NULL) || (need_field && rq->d_field == NULL))) {
    NoAllocSay(errbuf, errlen, "b200p_parse_device_noalloc: an output array this format writes is NULL");
    return B200P_E_INVALID;
  }
  out->index_bytes = rq->index_bytes;
  out->on_device = 1;
  out->text_bytes = rq->nbytes;
  if (rq->nbytes == 0) return B200P_OK;

#define NOALLOC_TRY(expr)                                                                     \
  do {                                                                                        \
    cudaError_t e_ = (expr);                                                                  \
    if (e_ != cudaSuccess) {                                                                  \
      if (errbuf != NULL && errlen != 0)                                                      \
        std::snprintf(errbuf, errlen, "b200p_parse_device_noalloc: CUDA error %s at line %d: %s", \
                      cudaGetErrorName(e_), __LINE__, cudaGetErrorString(e_));                \
      return B200P_E_CUDA;                                                                    \
    }                                                                                         \
  } while (0)

  DeviceScope on_device(rq->device);
  NOALLOC_TRY(on_device.status());
  NoAllocScratch *scratch = NULL;
  NOALLOC_TRY(NoAllocScratchFor(rq->device, &scratch));
  cudaStream_t stream = static_cast<cudaStream_t>(rq->stream);

  WalkParams p;
  std::memset(&p, 0, sizeof(p));
  p.text = static_cast<const uint8_t *>(rq->d_text);
  p.size = static_cast<uint32_t>(rq->nbytes);
  p.range_bytes = kNoAllocRangeBytes;
  p.num_ranges = static_cast<uint32_t>((rq->nbytes + kNoAllocRangeBytes - 1) / kNoAllocRangeBytes);
  p.label_column = rq->label_column;
  p.four = 4u;
  p.sixteen = 16u;
  p.counts = scratch->counts;
  p.num_tiles = (p.num_ranges + kScanTile - 1) / kScanTile;
  p.tile_base = scratch->tile_base;
  p.summary = &scratch->summary;
  p.offset = reinterpret_cast<unsigned long long *>(rq->d_offset);
  p.label = rq->d_label;
  p.weight = rq->d_weight;
  p.qid = reinterpret_cast<unsigned long long *>(rq->d_qid);
  p.field = rq->d_field;
  p.index = rq->d_index;
  p.value = rq->d_value;
  p.cap_rows = rq->cap_rows;
  p.cap_nnz = rq->cap_nnz;
  const bool wide = rq->index_bytes == 8;

  /* the summary (totals + error word) must start from zero; the counters need no clearing
     because pass 1 stores all five counters of every range it is launched for */
  NOALLOC_TRY(cudaMemsetAsync(&scratch->summary, 0, sizeof(ChunkSummary), stream));
  LaunchWalkPlain(rq->format, wide, p, stream, false);
  ScanTiles<<<p.num_tiles, kScanTile, 0, stream>>>(p.counts, p.num_ranges, p.tile_base, p.num_tiles);
  ScanFinish<<<1, kScanTile, 0, stream>>>(p.tile_base, p.num_tiles, p.num_ranges, &scratch->summary, NULL);
  if (rq->format == B200P_FORMAT_CSV) {
    ZeroLabels<<<296, 256, 0, stream>>>(p.label, &scratch->summary, p.cap_rows, p.cap_nnz);
  }
  LaunchWalkPlain(rq->format, wide, p, stream, true);
  NOALLOC_TRY(cudaGetLastError());
  ChunkSummary sum; /* on the caller's stack: nothing is allocated for the read-back either */
  NOALLOC_TRY(cudaMemcpyAsync(&sum, &scratch->summary, sizeof(sum), cudaMemcpyDeviceToHost, stream));
  NOALLOC_TRY(cudaStreamSynchronize(stream));
#undef NOALLOC_TRY

  out->rows = sum.total[kRows];
  out->nnz = sum.total[kNnz];
  out->n_weight = sum.total[kWeights];
  out->n_qid = sum.total[kQids];
  out->n_value = sum.total[kValues];
  out->n_field = need_field ? out->nnz : 0;
  if (sum.total[kRows] > rq->cap_rows || sum.total[kNnz] > rq->cap_nnz) {
    /* pass 2 wrote nothing; rows / nnz tell the caller how much room this chunk needs */
    if (errbuf != NULL && errlen != 0) {
      std::snprintf(errbuf, errlen,
                    "b200p_parse_device_noalloc: output arrays too small, chunk has %llu rows and "
                    "%llu entries", sum.total[kRows], sum.total[kNnz]);
    }
    return B200P_E_CAPACITY;
  }
  if (sum.error & kErrInternal) {
    NoAllocSay(errbuf, errlen, "internal consistency check failed (checked build): a queue or output index left its bounds");
    return B200P_E_CUDA;
  }
  if (sum.error & kErrNegativeIndex) {
    NoAllocSay(errbuf, errlen, "Check failed: sign == true (negative value in an unsigned index/field)");
    return B200P_E_PARSE;
  }
  if (sum.error & kErrValueCount) {
    NoAllocSay(errbuf, errlen, "Check failed: offset.back() == value.size() || value.size() == 0");
    return B200P_E_PARSE;
  }
  if (out->rows != 0) {
    out->offset = rq->d_offset;
    out->label = rq->d_label;
    out->weight = out->n_weight ? rq->d_weight : NULL;
    out->qid = out->n_qid ? rq->d_qid : NULL;
    out->field = out->n_field ? rq->d_field : NULL;
    out->index = out->nnz ? rq->d_index : NULL;
    out->value = out->n_value ? rq->d_value : NULL;
  }
  return B200P_OK;
}

}  // extern "C"
