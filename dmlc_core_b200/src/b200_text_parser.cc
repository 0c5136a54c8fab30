/*
 * dmlc::Parser<IndexType> on the B200: the drop-in for what the reference builds in
 * src/data.cc:21-85 (factories, CreateParser_, Parser<I>::Create) on top of
 * TextParserBase / LibSVMParser / LibFMParser / CSVParser / ThreadedParser.
 *
 * Host side only: a feeder thread reads whole-line chunks of the InputSplit partition straight
 * into the pinned staging buffers of the C ABI (include/dmlc_b200_parser.h) and submits them,
 * keeping `depth` chunks in flight on the GPU -- the role ThreadedInputSplit + ThreadedParser
 * play in the reference (src/io/threaded_input_split.h:29-79, src/data/parser.h:70-126) -- while
 * the consumer thread takes each parsed chunk as one RowBlock whose arrays live in pinned host
 * memory.  All parsing happens in the sm_100a kernels; there is no CPU fallback.
 */
#include <dmlc/data.h>
#include <dmlc/io.h>
#include <dmlc/logging.h>
#include <dmlc/parameter.h>
#include <dmlc/timer.h>

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <functional>
#include <limits>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

#include <sys/mman.h>

#include "../../include/dmlc_b200_parser.h"
#include "../csrc/host_pool.h"
#include "./uri_spec.h"
/*
 * DMLC_B200 is defined by this package's own <dmlc/base.h>.  Built against the reference's
 * headers instead (the true drop-in build) the parser takes whatever
 * InputSplit the reference's InputSplit::Create returns and copies each chunk into pinned memory;
 * built here it recognises its own TextSplit and lets it read straight into the pinned buffers.
 */
#if defined(DMLC_B200)
#include "./text_split.h"
#define B200_HAVE_TEXT_SPLIT 1
#else
#define B200_HAVE_TEXT_SPLIT 0
#endif

namespace dmlc {
namespace b200 {

static int EnvInt(const char *name, int fallback) {
  const char *v = std::getenv(name);
  return (v != NULL && *v != '\0') ? std::atoi(v) : fallback;
}

/* Which GPU a parser runs on.
   DMLC_B200_DEVICE = a CUDA ordinal, or "auto" = part_index modulo DMLC_B200_NUM_DEVICES
   (default 8), for a single process that opens one parser per part and wants them spread over
   the GPUs of the box.  Without the variable: the device that is current on the calling thread
   (a host program that called cudaSetDevice(k) gets its blocks parsed on GPU k); a process that
   has not touched CUDA yet but was started by a launcher gets LOCAL_RANK. */
static int DeviceFor(unsigned part_index) {
  const char *v = std::getenv("DMLC_B200_DEVICE");
  if (v != NULL && std::strcmp(v, "auto") == 0) {
    const int n = std::max(1, EnvInt("DMLC_B200_NUM_DEVICES", 8));
    return static_cast<int>(part_index % static_cast<unsigned>(n));
  }
  if (v != NULL && *v != '\0') return std::atoi(v);
  const int current = b200p_current_device();
  if (current > 0) return current;
  const int local = EnvInt("LOCAL_RANK", -1);
  const int count = b200p_device_count();
  if (local >= 0 && count > 0) return local % count;
  return std::max(0, current);
}

/*
 * One parser instance = one partition of the input on one GPU.
 * Environment knobs (GPU-side settings cannot ride in the URI: CSVParserParam::Init rejects
 * unknown keys, reference csv_parser.h:50 + parameter.h:405-418):
 *   DMLC_B200_DEVICE    CUDA device ordinal (default 0), or "auto" = part_index % DMLC_B200_NUM_DEVICES
 *   DMLC_B200_CHUNK_MB  chunk size in MiB (default 64, less for partitions under 384 MiB; the
 *                       reference uses 8, input_split_base.h:40)
 *   DMLC_B200_DEPTH     chunks in flight (default 3)
 *   DMLC_B200_READ_THREADS  threads of one parallel pread (default: 3/4 of the cores, at most 12,
 *                           divided by LOCAL_WORLD_SIZE)
 */
template <typename IndexType>
class B200TextParser : public Parser<IndexType> {
 public:
  B200TextParser(InputSplit *source, int format, int label_column, unsigned part_index = 0)
      : source_(source), text_(NULL), handle_(NULL),
        bytes_read_(0), exhausted_(false), feeder_running_(false), feeder_stop_(false),
        feeder_done_(false) {
    std::memset(&block_, 0, sizeof(block_));
#if B200_HAVE_TEXT_SPLIT
    text_ = dynamic_cast<LineSource *>(source);   /* TextSplit or StdinSplit */
#endif
    b200p_config cfg;
    std::memset(&cfg, 0, sizeof(cfg));
    cfg.format = format;
    cfg.index_bytes = static_cast<int32_t>(sizeof(IndexType));
    cfg.label_column = label_column;
    cfg.device = DeviceFor(part_index);
    chunk_bytes_ = static_cast<size_t>(std::max(1, EnvInt("DMLC_B200_CHUNK_MB", 64))) << 20;
    chunk_bytes_ = std::min<size_t>(chunk_bytes_, B200P_MAX_CHUNK_BYTES);
    if (EnvInt("DMLC_B200_CHUNK_MB", 0) <= 0 && PartitionBytes() != 0) {
      /* a small partition gets small chunks: the pinned staging and result buffers of the
         pipeline slots are the start-up cost, and six chunks are enough to overlap read, copy
         and parse (100 MB of LibSVM: first pass 150 ms with 64 MiB chunks, 3x less this way) */
      const size_t sixth = (PartitionBytes() / 6 + (1u << 20)) >> 20 << 20;
      chunk_bytes_ = std::min(chunk_bytes_, std::max<size_t>(sixth, 4u << 20));
    }
    cfg.chunk_bytes = chunk_bytes_;
    depth_ = std::max(1, std::min(8, EnvInt("DMLC_B200_DEPTH", 3)));
    cfg.depth = depth_;
    const int rc = b200p_create(&cfg, &handle_);
    if (rc != B200P_OK) {
      const std::string why = b200p_last_error(NULL);
      delete source_;
      source_ = NULL;
      LOG(FATAL) << "dmlc_core_b200: cannot create the GPU parser: " << why;
    }
    if (text_ == NULL) source_->HintChunkSize(chunk_bytes_);
  }

  virtual ~B200TextParser() {
    StopFeeder();
    b200p_destroy(handle_);
    delete source_;
  }

  virtual void BeforeFirst(void) {
    /* discard what is still in flight, then rewind the source (reference text_parser.h:40-42;
       like there, BytesRead keeps counting across passes) */
    StopFeeder();
    feeder_error_ = nullptr;
    b200p_rowblock dropped;
    while (b200p_pending(handle_) > 0) {
      /* b200p_next takes the chunk off the queue before it looks at it, so a chunk that fails
         to parse is gone all the same: keep draining, or the blocks behind it would come out
         at the head of the next pass.  Only a CUDA failure stops the loop. */
      const int rc = b200p_next(handle_, &dropped);
      if (rc == B200P_E_CUDA || rc == B200P_E_INVALID) break;
    }
    source_->BeforeFirst();
    exhausted_ = false;
  }

  virtual bool Next(void) {
    for (;;) {
      if (text_ != NULL) {
        if (!WaitForFeeder()) return false;
      } else {
        Feed();
        if (b200p_pending(handle_) == 0) return false;
      }
      b200p_rowblock b;
      const int rc = b200p_next(handle_, &b);
      if (text_ != NULL) {   /* a pipeline slot came free */
        { std::lock_guard<std::mutex> g(feeder_mu_); }
        feeder_cv_.notify_all();
      }
      Check(rc);
      bytes_read_ += b.text_bytes;
      if (b.rows == 0) continue;  /* chunks without rows are skipped (reference parser.h:33-38) */
      block_.size = static_cast<size_t>(b.rows);
      block_.offset = reinterpret_cast<const size_t *>(b.offset);
      block_.label = b.label;
      block_.weight = b.weight;
      block_.qid = b.qid;
      block_.field = static_cast<const IndexType *>(b.field);
      block_.index = static_cast<const IndexType *>(b.index);
      block_.value = b.value;
      return true;
    }
  }

  virtual const RowBlock<IndexType> &Value(void) const { return block_; }
  virtual size_t BytesRead(void) const { return bytes_read_; }
  /* largest feature index of the current block, reduced on the GPU from the array still
     resident there (b200p_block_max) */
  void BlockMax(uint64_t *max_index, uint64_t *max_field) {
    Check(b200p_block_max(handle_, max_index, max_field));
  }
  uint64_t BlockMaxIndex() {
    uint64_t mi = 0, mf = 0;
    BlockMax(&mi, &mf);
    return mi;
  }
  /* Next() with the block written straight behind what the caller already holds
     (b200p_next_to).  kRefused: the next block does not fit `dst` or needs an array `dst` lacks;
     it stays queued, *blk says what it holds, and Next() can still fetch it the ordinary way. */
  enum IntoResult { kIntoEnd, kIntoDone, kIntoRefused };
  IntoResult NextInto(const b200p_destination &dst, b200p_rowblock *blk) {
    for (;;) {
      if (text_ != NULL) {
        if (!WaitForFeeder()) return kIntoEnd;
      } else {
        Feed();
        if (b200p_pending(handle_) == 0) return kIntoEnd;
      }
      const int rc = b200p_next_to(handle_, &dst, blk);
      if ((rc == B200P_E_CAPACITY || rc == B200P_E_INVALID) && blk->text_bytes != 0) return kIntoRefused;
      if (text_ != NULL) {   /* a pipeline slot came free */
        { std::lock_guard<std::mutex> g(feeder_mu_); }
        feeder_cv_.notify_all();
      }
      Check(rc);
      bytes_read_ += blk->text_bytes;
      if (blk->rows == 0) continue;
      return kIntoDone;
    }
  }
  /* blocks will be asked for with NextInto: their arrays travel when the destination is known */
  void DeliverOnDemand() { b200p_set_delivery(handle_, B200P_DELIVER_ON_DEMAND); }
  /* pinned host memory that takes device-to-host copies at link speed (b200p_host_alloc) */
  void *HostAlloc(size_t bytes) {   /* NULL when the host will not pin that much */
    void *p = NULL;
    return b200p_host_alloc(handle_, bytes, &p) == B200P_OK ? p : NULL;
  }
  /* text bytes of this parser's partition when the source can tell (files), else 0 */
  size_t PartitionBytes() const {
#if B200_HAVE_TEXT_SPLIT
    if (text_ != NULL) return static_cast<size_t>(text_->PassBytes());
#endif
    return 0;
  }

 private:
  /*
   * Sources that can fill pinned memory themselves (TextSplit, StdinSplit).  The feeder thread loops { borrow the pinned staging buffer of a free
   * pipeline slot; fill it with the next whole lines of the partition (parallel pread);
   * b200p_submit } whenever fewer than `depth` chunks are in flight; the consumer thread only
   * calls b200p_next.  The C ABI allows exactly this pairing of threads on one handle.  An
   * exception in the feeder (I/O error) is kept and rethrown by Next() once the blocks submitted
   * before it have been delivered, like ThreadedIter does (include/dmlc/threadediter.h:456-466).
   */
  void FeederLoop() {
#if B200_HAVE_TEXT_SPLIT
    try {
      for (;;) {
        {
          std::unique_lock<std::mutex> g(feeder_mu_);
          feeder_cv_.wait(g, [this] { return feeder_stop_ || b200p_pending(handle_) < depth_; });
          if (feeder_stop_) break;
        }
        void *pinned = NULL;
        uint64_t cap = 0;
        const int rc = b200p_input_buffer(handle_, &pinned, &cap);
        if (rc == B200P_E_AGAIN) {   /* every slot taken: wait for the consumer to return one */
          std::unique_lock<std::mutex> g(feeder_mu_);
          feeder_cv_.wait_for(g, std::chrono::milliseconds(1));
          continue;
        }
        Check(rc);
        size_t n = text_->ReadLines(pinned, static_cast<size_t>(cap));
        const void *chunk = pinned;
        if (n == LineSource::kLineTooLong) {
          /* a single line longer than the staging buffer: read it through a growing heap
             buffer; b200p_submit then enlarges the slot */
          size_t big = static_cast<size_t>(cap);
          do {
            big *= 2;
            CHECK(big <= B200P_MAX_CHUNK_BYTES) << "a single line exceeds " << B200P_MAX_CHUNK_BYTES << " bytes";
            spill_.resize(big);
            n = text_->ReadLines(spill_.data(), big);
          } while (n == LineSource::kLineTooLong);
          chunk = spill_.data();
        }
        /* at the end of the partition an empty chunk hands the borrowed buffer back (it comes
           out of b200p_next as a block without rows, which Next() skips) */
        Check(b200p_submit(handle_, n == 0 ? pinned : chunk, n));
        { std::lock_guard<std::mutex> g(feeder_mu_); }
        feeder_cv_.notify_all();
        if (n == 0) break;
      }
    } catch (...) {
      feeder_error_ = std::current_exception();
    }
#endif
    {
      std::lock_guard<std::mutex> g(feeder_mu_);
      feeder_done_ = true;
    }
    feeder_cv_.notify_all();
  }

  /* consumer side: true when a chunk is in flight, false at the end of the partition */
  bool WaitForFeeder() {
    if (!feeder_running_ && !exhausted_) {
      feeder_stop_ = false;
      feeder_done_ = false;
      feeder_running_ = true;
      feeder_ = std::thread([this] { FeederLoop(); });
    }
    if (feeder_running_) {
      std::unique_lock<std::mutex> g(feeder_mu_);
      feeder_cv_.wait(g, [this] { return feeder_done_ || b200p_pending(handle_) > 0; });
    }
    if (b200p_pending(handle_) > 0) return true;
    /* the feeder has finished and everything it submitted has been delivered */
    StopFeeder();
    exhausted_ = true;
    if (feeder_error_) {
      std::exception_ptr e = feeder_error_;
      feeder_error_ = nullptr;
      std::rethrow_exception(e);
    }
    return false;
  }

  void StopFeeder() {
    if (!feeder_running_) return;
    {
      std::lock_guard<std::mutex> g(feeder_mu_);
      feeder_stop_ = true;
    }
    feeder_cv_.notify_all();
    feeder_.join();
    feeder_running_ = false;
  }

  /* any other InputSplit (the build against the reference's own headers): chunks are copied into
     the staging buffer by b200p_submit on the consumer thread (the Blob dies at the next
     NextChunk); oversized chunks are cut at line ends */
  void Feed() {
    while ((!exhausted_ || !backlog_.empty()) && b200p_pending(handle_) < depth_) {
      if (!backlog_.empty()) {
        Check(b200p_submit(handle_, backlog_.front().data(), backlog_.front().size()));
        backlog_.erase(backlog_.begin());
        continue;
      }
      InputSplit::Blob chunk;
      if (!source_->NextChunk(&chunk)) { exhausted_ = true; break; }
      const char *base = static_cast<const char *>(chunk.dptr);
      if (chunk.size <= B200P_MAX_CHUNK_BYTES) {
        Check(b200p_submit(handle_, base, chunk.size));
        continue;
      }
      size_t done = 0;
      while (done < chunk.size) {
        size_t n = std::min<size_t>(chunk.size - done, B200P_MAX_CHUNK_BYTES);
        if (done + n < chunk.size) {
          while (n > 1 && base[done + n - 1] != '\n' && base[done + n - 1] != '\r') --n;
        }
        backlog_.push_back(std::string(base + done, n));
        done += n;
      }
    }
  }

  void Check(int rc) {
    if (rc == B200P_OK) return;
    /* B200P_E_PARSE carries the reference's CHECK text; everything surfaces as dmlc::Error */
    LOG(FATAL) << b200p_last_error(handle_);
  }

  InputSplit *source_;
#if B200_HAVE_TEXT_SPLIT
  LineSource *text_;   /* the source again, when it can fill pinned memory itself */
#else
  void *text_;
#endif
  b200p_handle handle_;
  RowBlock<IndexType> block_;
  size_t bytes_read_;
  size_t chunk_bytes_;
  int depth_;
  bool exhausted_;
  std::vector<char> spill_;
  std::vector<std::string> backlog_;
  /* feeder thread (TextSplit sources) */
  std::thread feeder_;
  std::mutex feeder_mu_;
  std::condition_variable feeder_cv_;
  bool feeder_running_;   /* consumer thread only */
  bool feeder_stop_, feeder_done_;   /* under feeder_mu_ */
  std::exception_ptr feeder_error_;  /* written by the feeder before feeder_done_, read after */
};

/* ---- factories: same signatures and argument handling as reference src/data.cc:21-60 ---- */
template <typename IndexType>
Parser<IndexType> *CreateLibSVMParser(const std::string &path,
                                      const std::map<std::string, std::string> &args,
                                      unsigned part_index, unsigned num_parts) {
  InputSplit *source = InputSplit::Create(path.c_str(), part_index, num_parts, "text");
  return new B200TextParser<IndexType>(source, B200P_FORMAT_LIBSVM, -1, part_index);
}

template <typename IndexType>
Parser<IndexType> *CreateLibFMParser(const std::string &path,
                                     const std::map<std::string, std::string> &args,
                                     unsigned part_index, unsigned num_parts) {
  InputSplit *source = InputSplit::Create(path.c_str(), part_index, num_parts, "text");
  return new B200TextParser<IndexType>(source, B200P_FORMAT_LIBFM, -1, part_index);
}

template <typename IndexType>
Parser<IndexType> *CreateCSVParser(const std::string &path,
                                   const std::map<std::string, std::string> &args,
                                   unsigned part_index, unsigned num_parts) {
  /* CSVParserParam (reference csv_parser.h:22-32): format (must be "csv"), label_column
     (default -1); any other key is a ParamError (parameter.h:136-142,405-418) */
  int label_column = -1;
  for (std::map<std::string, std::string>::const_iterator it = args.begin(); it != args.end(); ++it) {
    if (it->first == "format") {
      CHECK_EQ(it->second, "csv");
    } else if (it->first == "label_column") {
      char *end = NULL;
      const long v = std::strtol(it->second.c_str(), &end, 10);
      if (end == it->second.c_str() || *end != '\0') {
        throw ParamError("Invalid Parameter format for label_column expect int but value='" +
                         it->second + "'");
      }
      label_column = static_cast<int>(v);
    } else {
      throw ParamError("Cannot find argument '" + it->first +
                       "', Possible Arguments:\n----------------\nformat : string, optional, "
                       "default='csv'\n    File format.\nlabel_column : int, optional, default=-1\n"
                       "    Column index that will put into label.\n");
    }
  }
  InputSplit *source = InputSplit::Create(path.c_str(), part_index, num_parts, "text");
  return new B200TextParser<IndexType>(source, B200P_FORMAT_CSV, label_column, part_index);
}

/* type lookup: reference CreateParser_, src/data.cc:62-85 */
template <typename IndexType>
Parser<IndexType> *CreateParserImpl(const char *uri, unsigned part_index, unsigned num_parts,
                                    const char *type) {
  std::string ptype = type;
  UriSpec spec(uri, part_index, num_parts);
  if (ptype == "auto") {
    ptype = spec.args.count("format") != 0 ? spec.args.at("format") : std::string("libsvm");
  }
  const ParserFactoryReg<IndexType> *entry = Registry<ParserFactoryReg<IndexType> >::Get()->Find(ptype);
  if (entry == NULL) LOG(FATAL) << "Unknown data type " << ptype;
  return (*entry->body)(spec.uri, spec.args, part_index, num_parts);
}

/*
 * Growable array for the in-memory iterator: 2 MiB-aligned and advised for transparent huge
 * pages (a multi-GB std::vector spends most of its append time in 4 KiB first-touch faults),
 * grown geometrically unless Reserve was told the final size.
 */
/* bytes one in-memory load may keep pinned (see InMemoryRowIter::Init) */
static size_t PinBudgetBytes() {
  const int mb = EnvInt("DMLC_B200_ITER_PIN_MB", -1);
  if (mb >= 0) return static_cast<size_t>(mb) << 20;
  size_t avail_kb = 0;
  if (std::FILE *f = std::fopen("/proc/meminfo", "r")) {
    char line[256];
    while (std::fgets(line, sizeof(line), f) != NULL) {
      unsigned long long kb = 0;
      if (std::sscanf(line, "MemAvailable: %llu kB", &kb) == 1) {
        avail_kb = static_cast<size_t>(kb);
        break;
      }
    }
    std::fclose(f);
  }
  if (avail_kb == 0) return static_cast<size_t>(8) << 30;
  const int ranks = std::max(1, EnvInt("LOCAL_WORLD_SIZE", 1));
  return (avail_kb << 10) / 4 / static_cast<size_t>(ranks);
}

template <typename T>
class HugeArray {
 public:
  HugeArray() : p_(NULL), n_(0), cap_(0), pinned_(false) {}
  ~HugeArray() { Release(); }
  /* from now on memory comes from `alloc` (pinned by the GPU library, freed with b200p_host_free) */
  void UsePinned(const std::function<void *(size_t)> &alloc) { alloc_ = alloc; }
  T *End() { return p_ + n_; }
  size_t Room() const { return cap_ - n_; }
  void Commit(size_t k) { n_ += k; }
  size_t size() const { return n_; }
  bool empty() const { return n_ == 0; }
  const T *data() const { return n_ != 0 ? p_ : NULL; }   /* NULL when empty, like BeginPtr */
  const T &back() const { return p_[n_ - 1]; }
  void Reserve(size_t cap) {
    if (cap <= cap_) return;
    const size_t kAlign = 2u << 20;
    const size_t bytes = (cap * sizeof(T) + kAlign - 1) / kAlign * kAlign;
    void *q = NULL;
    bool pinned = false;
    if (alloc_) {
      q = alloc_(bytes);
      pinned = q != NULL;   /* a host that refuses to pin this much still gets ordinary memory */
    }
    if (q == NULL) {
      CHECK(posix_memalign(&q, kAlign, bytes) == 0) << "out of memory growing a RowBlock array";
#ifdef MADV_HUGEPAGE
      madvise(q, bytes, MADV_HUGEPAGE);
#endif
    }
    if (n_ != 0) std::memcpy(q, p_, n_ * sizeof(T));
    Release();
    pinned_ = pinned;
    p_ = static_cast<T *>(q);
    cap_ = bytes / sizeof(T);
  }
  /* room for k more elements; returns where they go */
  T *Extend(size_t k) {
    if (n_ + k > cap_) Reserve(std::max(n_ + k, cap_ + cap_ / 2));
    T *at = p_ + n_;
    n_ += k;
    return at;
  }
  void push_back(const T &v) { *Extend(1) = v; }

 private:
  void Release() {
    if (pinned_) b200p_host_free(p_); else std::free(p_);
    p_ = NULL;
  }
  T *p_;
  size_t n_, cap_;
  bool pinned_;
  std::function<void *(size_t)> alloc_;
  HugeArray(const HugeArray &);
  void operator=(const HugeArray &);
};

/*
 * RowBlockIter back-end: everything in host memory (reference BasicRowIter,
 * src/data/basic_row_iter.h:62-82, with RowBlockContainer::Push, row_block.h:122-166).
 * The parser's feeder thread and the GPU keep running while this thread appends a block, so the
 * load is bound by how fast blocks are copied: the arrays are sized from the first block's
 * density and the partition's byte length (one allocation, huge pages), and a block is copied
 * in by a small pool.
 */
template <typename IndexType>
class InMemoryRowIter : public RowBlockIter<IndexType> {
 public:
  explicit InMemoryRowIter(Parser<IndexType> *parser) : at_head_(true), max_index_(0), pool_(NULL) {
    offset_.push_back(0);
    const double t0 = GetTime();
    /* the B200 parser reduces each block's largest index on the GPU, where the array still
       sits; any other parser gets the reference's per-element scan */
    B200TextParser<IndexType> *gpu = dynamic_cast<B200TextParser<IndexType> *>(parser);
    const size_t total_bytes = gpu != NULL ? gpu->PartitionBytes() : 0;
    try {
      bool first = true;
      double t_next = 0, t_append = 0, t_max = 0;   /* DMLC_B200_ITER_TRACE: where the load went */
      const bool direct = gpu != NULL && std::getenv("DMLC_B200_ITER_COPY") == NULL;
      if (direct) gpu->DeliverOnDemand();
      for (;;) {
        double ta = GetTime();
        if (direct && !first) {
          /* every block after the first lands behind what is already here, no block in between */
          b200p_destination dst;
          b200p_rowblock blk;
          AimAtEnds(&dst);
          const typename B200TextParser<IndexType>::IntoResult got = gpu->NextInto(dst, &blk);
          if (got == B200TextParser<IndexType>::kIntoEnd) break;
          if (got == B200TextParser<IndexType>::kIntoDone) {
            t_next += GetTime() - ta;
            ta = GetTime();
            CommitBlock(blk);
            max_index_ = std::max(max_index_, static_cast<IndexType>(gpu->BlockMaxIndex()));
            t_max += GetTime() - ta;
            continue;
          }
          if (MakeRoomFor(blk)) continue;   /* grown: ask again */
          /* the block needs an array this load never had (weights or qids appearing late):
             take it the ordinary way */
        }
        if (!parser->Next()) break;
        t_next += GetTime() - ta;
        ta = GetTime();
        const RowBlock<IndexType> &b = parser->Value();
        if (first && direct) {
          /* DMLC_B200_ITER_NO_PIN: behave like a host that refuses to pin (ordinary memory) */
          const bool may_pin = std::getenv("DMLC_B200_ITER_NO_PIN") == NULL;
          /* Pinned pages are unswappable and no rlimit caps them, so the arrays of one load pin
             at most DMLC_B200_ITER_PIN_MB (default: a quarter of MemAvailable, divided between
             the local ranks); what does not fit is ordinary memory -- the copies into it go
             through the driver's staging, slower, nothing else changes. */
          std::shared_ptr<size_t> budget(new size_t(PinBudgetBytes()));
          const std::function<void *(size_t)> pinned = [gpu, may_pin, budget](size_t bytes) {
            if (!may_pin || bytes > *budget) return static_cast<void *>(NULL);
            void *p = gpu->HostAlloc(bytes);
            if (p != NULL) *budget -= bytes;
            return p;
          };
          offset_.UsePinned(pinned);
          label_.UsePinned(pinned);
          weight_.UsePinned(pinned);
          qid_.UsePinned(pinned);
          field_.UsePinned(pinned);
          index_.UsePinned(pinned);
          value_.UsePinned(pinned);
        }
        if (first && total_bytes != 0 && parser->BytesRead() != 0) {
          /* size everything once: this block's rows and entries per text byte, plus 5 % */
          const double scale = 1.05 * static_cast<double>(total_bytes) / parser->BytesRead();
          const size_t rows = static_cast<size_t>(b.size * scale) + 1024;
          const size_t nnz = static_cast<size_t>((b.offset[b.size] - b.offset[0]) * scale) + 1024;
          offset_.Reserve(rows + 1);
          label_.Reserve(rows);
          if (b.weight != NULL) weight_.Reserve(rows);
          if (b.qid != NULL) qid_.Reserve(rows);
          if (b.field != NULL) field_.Reserve(nnz);
          index_.Reserve(nnz);
          if (b.value != NULL) value_.Reserve(nnz);
        }
        first = false;
        Append(b, gpu == NULL);
        t_append += GetTime() - ta;
        ta = GetTime();
        if (gpu != NULL) {
          max_index_ = std::max(max_index_, static_cast<IndexType>(gpu->BlockMaxIndex()));
        }
        t_max += GetTime() - ta;
      }
      if (std::getenv("DMLC_B200_ITER_TRACE") != NULL) {
        std::fprintf(stderr, "iter trace: waiting for blocks %.3f s, appending %.3f s, max index %.3f s\n",
                     t_next, t_append, t_max);
      }
    } catch (...) {
      delete pool_;
      delete parser;
      throw;
    }
    delete pool_;
    pool_ = NULL;
    const double dt = GetTime() - t0;
    LOG(INFO) << "finish reading at " << (parser->BytesRead() >> 20UL) / (dt > 0 ? dt : 1e-9) << " MB/sec";
    delete parser;
    block_.size = label_.size();
    block_.offset = offset_.data();
    block_.label = label_.data();
    block_.weight = weight_.data();
    block_.qid = qid_.data();
    block_.field = field_.data();
    block_.index = index_.data();
    block_.value = value_.data();
    /* same consistency rules as RowBlockContainer::GetBlock (row_block.h:173-177) */
    CHECK_EQ(offset_.back(), index_.size());
    CHECK(offset_.back() == value_.size() || value_.size() == 0);
  }
  virtual void BeforeFirst(void) { at_head_ = true; }
  virtual bool Next(void) {
    const bool had = at_head_;
    at_head_ = false;
    return had;
  }
  virtual const RowBlock<IndexType> &Value(void) const { return block_; }
  virtual size_t NumCol(void) const { return static_cast<size_t>(max_index_) + 1; }

 private:
  struct Piece {
    void *dst;
    const void *src;
    size_t bytes;
  };
  /* destination of the next block: the ends of the arrays, and how much room they have */
  void AimAtEnds(b200p_destination *dst) {
    std::memset(dst, 0, sizeof(*dst));
    dst->offset = reinterpret_cast<uint64_t *>(offset_.End() - 1);   /* rewrites the last entry in place */
    dst->label = label_.End();
    dst->weight = weight_.Room() != 0 ? weight_.End() : NULL;
    dst->qid = qid_.Room() != 0 ? qid_.End() : NULL;
    dst->field = field_.Room() != 0 ? field_.End() : NULL;
    dst->index = index_.End();
    dst->value = value_.Room() != 0 ? value_.End() : NULL;
    size_t rows = std::min(offset_.Room(), label_.Room());
    if (!weight_.empty()) rows = std::min(rows, weight_.Room());
    if (!qid_.empty()) rows = std::min(rows, qid_.Room());
    size_t nnz = index_.Room();
    if (!field_.empty()) nnz = std::min(nnz, field_.Room());
    if (!value_.empty()) nnz = std::min(nnz, value_.Room());
    dst->cap_rows = rows;
    dst->cap_nnz = nnz;
    dst->offset_base = offset_.back();
  }
  /* the block NextInto wrote is now part of the arrays (same lengths Append would add) */
  void CommitBlock(const b200p_rowblock &blk) {
    offset_.Commit(blk.rows);
    label_.Commit(blk.rows);
    if (blk.weight != NULL) weight_.Commit(blk.rows);
    if (blk.qid != NULL) qid_.Commit(blk.rows);
    if (blk.field != NULL) field_.Commit(blk.nnz);
    index_.Commit(blk.nnz);
    if (blk.value != NULL) value_.Commit(blk.nnz);
  }
  /* after a refusal: grow what the block needs; false when it needs an array that has never
     been used in this load (then the ordinary path appends it) */
  bool MakeRoomFor(const b200p_rowblock &blk) {
    if ((blk.n_weight != 0 && weight_.empty()) || (blk.n_qid != 0 && qid_.empty()) ||
        (blk.n_field != 0 && field_.empty()) || (blk.n_value != 0 && value_.empty())) {
      return false;
    }
    const size_t rows = static_cast<size_t>(blk.rows), nnz = static_cast<size_t>(blk.nnz);
    Grow(&offset_, rows);
    Grow(&label_, rows);
    if (!weight_.empty()) Grow(&weight_, rows);
    if (!qid_.empty()) Grow(&qid_, rows);
    if (!field_.empty()) Grow(&field_, nnz);
    Grow(&index_, nnz);
    if (!value_.empty()) Grow(&value_, nnz);
    return true;
  }
  template <typename T>
  static void Grow(HugeArray<T> *a, size_t more) {
    if (a->Room() < more) a->Reserve(a->size() + std::max(more, a->size() / 2));
  }
  template <typename T>
  void Plan(HugeArray<T> *to, const T *from, size_t count, std::vector<Piece> *plan) {
    if (from == NULL || count == 0) return;
    char *dst = reinterpret_cast<char *>(to->Extend(count));
    const char *src = reinterpret_cast<const char *>(from);
    const size_t kPiece = 4u << 20;
    for (size_t done = 0, bytes = count * sizeof(T); done < bytes; done += kPiece) {
      const Piece p = {dst + done, src + done, std::min(kPiece, bytes - done)};
      plan->push_back(p);
    }
  }
  void Append(const RowBlock<IndexType> &b, bool scan_max) {
    const size_t base = b.offset[0], nnz = b.offset[b.size] - base, shift = offset_.back();
    std::vector<Piece> plan;
    Plan(&label_, b.label, b.size, &plan);
    Plan(&weight_, b.weight, b.size, &plan);
    Plan(&qid_, b.qid, b.size, &plan);
    Plan(&field_, b.field != NULL ? b.field + base : NULL, nnz, &plan);
    Plan(&index_, b.index != NULL ? b.index + base : NULL, nnz, &plan);
    Plan(&value_, b.value != NULL ? b.value + base : NULL, nnz, &plan);
    size_t *rebased = offset_.Extend(b.size);
    if (plan.size() > 1) {
      if (pool_ == NULL) {
        const unsigned hw = std::thread::hardware_concurrency();
        pool_ = new b200p::HostPool(static_cast<int>(std::max(1u, std::min(4u, hw / 4u))));
      }
      pool_->Run(static_cast<int>(plan.size()), [&plan](int i) {
        std::memcpy(plan[i].dst, plan[i].src, plan[i].bytes);
      });
    } else if (plan.size() == 1) {
      std::memcpy(plan[0].dst, plan[0].src, plan[0].bytes);
    }
    for (size_t i = 1; i <= b.size; ++i) rebased[i - 1] = shift + b.offset[i] - base;
    if (scan_max) {
      for (size_t i = 0; i < nnz; ++i) max_index_ = std::max(max_index_, b.index[base + i]);
    }
  }
  bool at_head_;
  IndexType max_index_;
  RowBlock<IndexType> block_;
  b200p::HostPool *pool_;   /* only while loading */
  HugeArray<size_t> offset_;
  HugeArray<real_t> label_, weight_, value_;
  HugeArray<uint64_t> qid_;
  HugeArray<IndexType> field_, index_;
};

/*
 * RowBlockIter back-end with a disk cache of parsed pages ("uri#cachefile"): reference
 * DiskRowIter, src/data/disk_row_iter.h:27-141.  The page format is the reference's
 * RowBlockContainer::Save (src/data/row_block.h:189-201 over serializer.h:105-112): seven arrays,
 * each a u64 element count followed by the raw elements (offset u64, label f32, weight f32,
 * qid u64, field I, index I, value f32), then max_field and max_index as IndexType.  Caches
 * written here load in the reference and the other way round; a page is closed once it holds
 * >= 64 MiB of arrays (disk_row_iter.h:32,123).
 */
template <typename IndexType>
class CachedRowIter : public RowBlockIter<IndexType> {
 public:
  static const size_t kPageSize = 64UL << 20UL;

  /* true when `cache_file` exists and starts with a readable page */
  static bool Usable(const std::string &cache_file) {
    std::FILE *f = std::fopen(cache_file.c_str(), "rb");
    if (f == NULL) return false;
    uint64_t n = 0;
    const bool ok = std::fread(&n, sizeof(n), 1, f) == 1 && n >= 1;
    std::fclose(f);
    return ok;
  }

  /* parser == NULL: the cache must already exist */
  CachedRowIter(Parser<IndexType> *parser, const std::string &cache_file)
      : cache_file_(cache_file), fi_(NULL), num_col_(0), held_(-1), next_slot_(0), loading_(false),
        stop_(false), eof_(false) {
    ready_[0] = ready_[1] = false;
    if (parser != NULL) {
      try {
        Build(parser);
      } catch (...) {
        delete parser;
        throw;
      }
      delete parser;
    }
    fi_ = std::fopen(cache_file_.c_str(), "rb");
    CHECK(fi_ != NULL) << "failed to build cache file " << cache_file_;
    if (parser == NULL) {
      /* NumCol of an existing cache: the largest max_index over its pages */
      while (LoadPage(&pages_[0])) num_col_ = std::max(num_col_, static_cast<size_t>(pages_[0].max_index) + 1);
      std::fseek(fi_, 0, SEEK_SET);
    }
  }
  virtual ~CachedRowIter() {
    StopLoader();
    if (fi_ != NULL) std::fclose(fi_);
  }
  virtual void BeforeFirst(void) {
    StopLoader();
    std::fseek(fi_, 0, SEEK_SET);
    ready_[0] = ready_[1] = false;
    held_ = -1;
    next_slot_ = 0;
    eof_ = false;
  }
  /* Pages come through a loader thread with two page buffers, so that page k+1 is read from disk
     while the caller works on page k (the reference replays its cache through a ThreadedIter,
     src/data/disk_row_iter.h:96-108). */
  virtual bool Next(void) {
    std::unique_lock<std::mutex> lk(mu_);
    if (held_ >= 0) {          /* hand the previous page's buffer back to the loader */
      ready_[held_] = false;
      held_ = -1;
      cv_.notify_all();
    }
    if (!loading_ && !eof_) {
      stop_ = false;
      loading_ = true;
      loader_ = std::thread([this] { LoaderLoop(); });
    }
    const int slot = next_slot_;
    cv_.wait(lk, [this, slot] { return ready_[slot] || eof_; });
    if (!ready_[slot]) {
      if (loader_error_) {
        std::exception_ptr e = loader_error_;
        loader_error_ = nullptr;
        std::rethrow_exception(e);
      }
      return false;
    }
    held_ = slot;
    next_slot_ = 1 - slot;
    const Page &pg = pages_[slot];
    block_.size = pg.label.size();
    block_.offset = BeginPtr(pg.offset);
    block_.label = BeginPtr(pg.label);
    block_.weight = BeginPtr(pg.weight);
    block_.qid = BeginPtr(pg.qid);
    block_.field = BeginPtr(pg.field);
    block_.index = BeginPtr(pg.index);
    block_.value = BeginPtr(pg.value);
    return true;
  }
  virtual const RowBlock<IndexType> &Value(void) const { return block_; }
  virtual size_t NumCol(void) const { return num_col_; }

 private:
  struct Page {
    std::vector<size_t> offset;
    std::vector<real_t> label, weight, value;
    std::vector<uint64_t> qid;
    std::vector<IndexType> field, index;
    IndexType max_index, max_field;
    Page() : max_index(0), max_field(0) {}
    size_t MemCost() const {
      return offset.size() * sizeof(size_t) + (label.size() + weight.size() + value.size()) * sizeof(real_t) +
             qid.size() * sizeof(size_t) + (field.size() + index.size()) * sizeof(IndexType);
    }
    void Clear() {
      offset.assign(1, 0);
      label.clear(); weight.clear(); qid.clear(); field.clear(); index.clear(); value.clear();
      max_index = 0;
      max_field = 0;
    }
  };

  void LoaderLoop() {
    int slot = next_slot_;
    try {
      for (;;) {
        {
          std::unique_lock<std::mutex> lk(mu_);
          cv_.wait(lk, [this, slot] { return stop_ || !ready_[slot]; });
          if (stop_) break;
        }
        const bool got = LoadPage(&pages_[slot]);   /* the slot is free: nobody else touches it */
        std::lock_guard<std::mutex> lk(mu_);
        if (!got) {
          eof_ = true;
          cv_.notify_all();
          break;
        }
        ready_[slot] = true;
        cv_.notify_all();
        slot = 1 - slot;
      }
    } catch (...) {
      std::lock_guard<std::mutex> lk(mu_);
      loader_error_ = std::current_exception();
      eof_ = true;
      cv_.notify_all();
    }
  }
  void StopLoader() {
    {
      std::lock_guard<std::mutex> lk(mu_);
      if (!loading_) return;
      stop_ = true;
    }
    cv_.notify_all();
    loader_.join();
    loading_ = false;
  }

  /* one array of a page: element count, then the elements (reference include/dmlc/serializer.h:105-112) */
  template <typename T>
  static void WriteArray(std::FILE *f, const T *data, size_t count) {
    const uint64_t n = count;
    CHECK(std::fwrite(&n, sizeof(n), 1, f) == 1) << "cache write failed";
    if (n != 0) CHECK(std::fwrite(data, sizeof(T), count, f) == count) << "cache write failed";
  }
  template <typename T>
  static void WriteVec(std::FILE *f, const std::vector<T> &v) { WriteArray(f, BeginPtr(v), v.size()); }
  template <typename T>
  static bool ReadVec(std::FILE *f, std::vector<T> *v) {
    uint64_t n = 0;
    if (std::fread(&n, sizeof(n), 1, f) != 1) return false;
    v->resize(static_cast<size_t>(n));
    return n == 0 || std::fread(&(*v)[0], sizeof(T), v->size(), f) == v->size();
  }
  void WriteTail(std::FILE *fo, IndexType max_field, IndexType max_index) {
    CHECK(std::fwrite(&max_field, sizeof(IndexType), 1, fo) == 1) << "cache write failed";
    CHECK(std::fwrite(&max_index, sizeof(IndexType), 1, fo) == 1) << "cache write failed";
    num_col_ = std::max(num_col_, static_cast<size_t>(max_index) + 1);
  }
  void SavePage(std::FILE *fo, const Page &pg) {
    WriteVec(fo, pg.offset); WriteVec(fo, pg.label); WriteVec(fo, pg.weight); WriteVec(fo, pg.qid);
    WriteVec(fo, pg.field); WriteVec(fo, pg.index); WriteVec(fo, pg.value);
    WriteTail(fo, pg.max_field, pg.max_index);
  }
  /* a block that is a page by itself goes to the file straight from the parser's (pinned) arrays:
     same page layout (RowBlockContainer::Save, reference src/data/row_block.h:189-201), no copy */
  void SaveBlockAsPage(std::FILE *fo, const RowBlock<IndexType> &b, IndexType max_index, IndexType max_field) {
    const size_t base = b.offset[0], nnz = b.offset[b.size] - base;
    if (base == 0) {
      WriteArray(fo, b.offset, b.size + 1);
    } else {
      std::vector<size_t> rebased(b.size + 1);
      for (size_t i = 0; i <= b.size; ++i) rebased[i] = b.offset[i] - base;
      WriteVec(fo, rebased);
    }
    WriteArray(fo, b.label, b.size);
    WriteArray(fo, b.weight, b.weight != NULL ? b.size : 0);
    WriteArray(fo, b.qid, b.qid != NULL ? b.size : 0);
    WriteArray(fo, b.field != NULL ? b.field + base : b.field, b.field != NULL ? nnz : 0);
    WriteArray(fo, b.index != NULL ? b.index + base : b.index, b.index != NULL ? nnz : 0);
    WriteArray(fo, b.value != NULL ? b.value + base : b.value, b.value != NULL ? nnz : 0);
    WriteTail(fo, max_field, max_index);
  }
  bool LoadPage(Page *pg) {
    if (!ReadVec(fi_, &pg->offset)) return false;
    CHECK(ReadVec(fi_, &pg->label) && ReadVec(fi_, &pg->weight) && ReadVec(fi_, &pg->qid) &&
          ReadVec(fi_, &pg->field) && ReadVec(fi_, &pg->index) && ReadVec(fi_, &pg->value) &&
          std::fread(&pg->max_field, sizeof(IndexType), 1, fi_) == 1 &&
          std::fread(&pg->max_index, sizeof(IndexType), 1, fi_) == 1)
        << "Bad RowBlock format";
    return true;
  }
  static size_t BlockCost(const RowBlock<IndexType> &b) {
    const size_t nnz = b.offset[b.size] - b.offset[0];
    return (b.size + 1) * sizeof(size_t) + b.size * sizeof(real_t) * (b.weight != NULL ? 2 : 1) +
           (b.qid != NULL ? b.size * sizeof(uint64_t) : 0) +
           nnz * (sizeof(IndexType) * (b.field != NULL ? 2 : 1) + (b.value != NULL ? sizeof(real_t) : 0));
  }
  void Build(Parser<IndexType> *parser) {
    std::FILE *fo = std::fopen(cache_file_.c_str(), "wb");
    CHECK(fo != NULL) << "cannot write cache file " << cache_file_;
    B200TextParser<IndexType> *gpu = dynamic_cast<B200TextParser<IndexType> *>(parser);
    const double t0 = GetTime();
    Page &acc = pages_[0];      /* small blocks are gathered here until they make a page */
    acc.Clear();
    try {
      while (parser->Next()) {
        const RowBlock<IndexType> &b = parser->Value();
        const size_t base = b.offset[0], nnz = b.offset[b.size] - base;
        IndexType max_index = 0, max_field = 0;
        if (gpu != NULL) {
          uint64_t mi = 0, mf = 0;
          gpu->BlockMax(&mi, &mf);
          max_index = static_cast<IndexType>(mi);
          max_field = static_cast<IndexType>(mf);
        } else {
          for (size_t i = 0; i < nnz; ++i) max_index = std::max(max_index, b.index[base + i]);
          if (b.field != NULL) {
            for (size_t i = 0; i < nnz; ++i) max_field = std::max(max_field, b.field[base + i]);
          }
        }
        if (acc.label.empty() && BlockCost(b) >= kPageSize / 2) {
          SaveBlockAsPage(fo, b, max_index, max_field);
          continue;
        }
        const size_t shift = acc.offset.back();
        acc.label.insert(acc.label.end(), b.label, b.label + b.size);
        if (b.weight != NULL) acc.weight.insert(acc.weight.end(), b.weight, b.weight + b.size);
        if (b.qid != NULL) acc.qid.insert(acc.qid.end(), b.qid, b.qid + b.size);
        if (b.field != NULL) acc.field.insert(acc.field.end(), b.field + base, b.field + base + nnz);
        acc.index.insert(acc.index.end(), b.index + base, b.index + base + nnz);
        if (b.value != NULL) acc.value.insert(acc.value.end(), b.value + base, b.value + base + nnz);
        for (size_t i = 1; i <= b.size; ++i) acc.offset.push_back(shift + b.offset[i] - base);
        acc.max_index = std::max(acc.max_index, max_index);
        acc.max_field = std::max(acc.max_field, max_field);
        if (acc.MemCost() >= kPageSize) {
          SavePage(fo, acc);
          acc.Clear();
        }
      }
      if (!acc.label.empty()) SavePage(fo, acc);
    } catch (...) {
      std::fclose(fo);
      std::remove(cache_file_.c_str());
      throw;
    }
    std::fclose(fo);
    acc.Clear();
    const double dt = GetTime() - t0;
    LOG(INFO) << "finish reading at " << (parser->BytesRead() >> 20UL) / (dt > 0 ? dt : 1e-9) << " MB/sec";
  }

  std::string cache_file_;
  std::FILE *fi_;
  size_t num_col_;
  RowBlock<IndexType> block_;
  Page pages_[2];
  /* loader thread state, under mu_ */
  std::mutex mu_;
  std::condition_variable cv_;
  std::thread loader_;
  bool ready_[2];
  int held_, next_slot_;
  bool loading_, stop_, eof_;
  std::exception_ptr loader_error_;
};

/* reference CreateIter_, src/data.cc:87-107: "#cachefile" selects the disk-backed iterator.
   Like there (:95-97), the parser is created from spec.uri -- the URI without its "?key=value"
   part -- so arguments such as format= or label_column= reach the parser only through
   Parser::Create, never through RowBlockIter::Create (with type "auto" the iterator therefore
   always parses LibSVM, and a CSV iterator has no label column). */
template <typename IndexType>
RowBlockIter<IndexType> *CreateIterImpl(const char *uri, unsigned part_index, unsigned num_parts,
                                        const char *type) {
  UriSpec spec(uri, part_index, num_parts);
  if (!spec.cache_file.empty()) {
    if (CachedRowIter<IndexType>::Usable(spec.cache_file)) {
      return new CachedRowIter<IndexType>(NULL, spec.cache_file);  /* later epochs skip parsing */
    }
    return new CachedRowIter<IndexType>(
        CreateParserImpl<IndexType>(spec.uri.c_str(), part_index, num_parts, type), spec.cache_file);
  }
  return new InMemoryRowIter<IndexType>(
      CreateParserImpl<IndexType>(spec.uri.c_str(), part_index, num_parts, type));
}

}  // namespace b200

/* ---- the public entry points, specialised for the two index widths like src/data.cc:113-147 */
template <>
Parser<uint32_t> *Parser<uint32_t>::Create(const char *uri, unsigned part_index, unsigned num_parts,
                                           const char *type) {
  return b200::CreateParserImpl<uint32_t>(uri, part_index, num_parts, type);
}
template <>
Parser<uint64_t> *Parser<uint64_t>::Create(const char *uri, unsigned part_index, unsigned num_parts,
                                           const char *type) {
  return b200::CreateParserImpl<uint64_t>(uri, part_index, num_parts, type);
}
template <>
RowBlockIter<uint32_t> *RowBlockIter<uint32_t>::Create(const char *uri, unsigned part_index,
                                                       unsigned num_parts, const char *type) {
  return b200::CreateIterImpl<uint32_t>(uri, part_index, num_parts, type);
}
template <>
RowBlockIter<uint64_t> *RowBlockIter<uint64_t>::Create(const char *uri, unsigned part_index,
                                                       unsigned num_parts, const char *type) {
  return b200::CreateIterImpl<uint64_t>(uri, part_index, num_parts, type);
}

/* registry singletons + the five stock registrations (src/data.cc:150-158; csv: uint32_t only) */
DMLC_REGISTRY_ENABLE(ParserFactoryReg<uint32_t>);
DMLC_REGISTRY_ENABLE(ParserFactoryReg<uint64_t>);
DMLC_REGISTER_DATA_PARSER(uint32_t, libsvm, b200::CreateLibSVMParser<uint32_t>);
DMLC_REGISTER_DATA_PARSER(uint64_t, libsvm, b200::CreateLibSVMParser<uint64_t>);
DMLC_REGISTER_DATA_PARSER(uint32_t, libfm, b200::CreateLibFMParser<uint32_t>);
DMLC_REGISTER_DATA_PARSER(uint64_t, libfm, b200::CreateLibFMParser<uint64_t>);
DMLC_REGISTER_DATA_PARSER(uint32_t, csv, b200::CreateCSVParser<uint32_t>);

}  // namespace dmlc
