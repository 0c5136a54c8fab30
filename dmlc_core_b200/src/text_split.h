/*
 * TextSplit: the "text" InputSplit for local files, built for feeding a GPU.
 *
 * Partition semantics are the reference's (InputSplitBase::Init/ResetPartition,
 * src/io/input_split_base.cc:13-64; LineSplitter::SeekRecordBegin / FindLastRecordBegin,
 * src/io/line_split.cc:9-34): the files behind the URI form one logical byte stream of length T;
 * part k of n is [ceil(T/n)*k, ceil(T/n)*(k+1)) with each end that is not a file start moved
 * forward past the next end-of-line run; chunks are cut after the last EOL that fits.
 *
 * What differs is the mechanics: files are read with positional reads (pread) fanned out over a
 * few threads, straight into a buffer the caller names -- the GPU parser passes the pinned
 * staging buffer of its next pipeline slot -- so there is no fread-into-vector plus overflow
 * memcpy (src/io/input_split_base.cc:211-258) and no prefetch thread
 * (src/io/threaded_input_split.h:29-42) between the page cache and the DMA engine.
 */
#ifndef DMLC_CORE_B200_TEXT_SPLIT_H_
#define DMLC_CORE_B200_TEXT_SPLIT_H_

#include <dmlc/io.h>

#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

namespace b200p {
class HostPool;
}

namespace dmlc {
namespace b200 {

/*
 * What the GPU parser's feeder thread needs from a source: whole lines written straight into
 * memory it names (the pinned staging buffer of a pipeline slot).
 */
class LineSource {
 public:
  static const size_t kLineTooLong = static_cast<size_t>(-1);
  virtual ~LineSource() {}
  /*
   * Fill `buf` (capacity `cap`) with the next run of whole lines.  Returns the number of bytes
   * (0 at the end) or kLineTooLong when not even one line fits in `cap` (retry with a larger
   * buffer; nothing is lost).
   */
  virtual size_t ReadLines(void *buf, size_t cap) = 0;
  /* bytes this source will deliver in one pass, 0 when it cannot know (a pipe) */
  virtual uint64_t PassBytes() const = 0;
};

class TextSplit : public InputSplit, public LineSource {
 public:

  TextSplit(const char *uri, unsigned part_index, unsigned num_parts);
  virtual ~TextSplit();

  virtual void HintChunkSize(size_t chunk_size);
  virtual size_t GetTotalSize(void);
  virtual void BeforeFirst(void);
  virtual bool NextRecord(Blob *out_rec);
  virtual bool NextChunk(Blob *out_chunk);
  virtual void ResetPartition(unsigned part_index, unsigned num_parts);

  virtual size_t ReadLines(void *buf, size_t cap);
  virtual uint64_t PassBytes() const { return end_ - begin_; }

  /* adjusted byte range of the partition within the logical stream */
  uint64_t PartitionBegin() const { return begin_; }
  uint64_t PartitionEnd() const { return end_; }
  /* threads used for one ReadLines call */
  void SetReadThreads(int n) { read_threads_ = n < 1 ? 1 : n; }

 private:
  struct File {
    std::string path;
    uint64_t size;
    int fd;
    const char *map;   /* read-only mapping of the file (DMLC_B200_READ_MMAP=1), else NULL */
  };
  void ExpandUri(const std::string &uri);
  void AddPath(const std::string &path, bool from_listing);
  size_t FileAt(uint64_t offset) const;            /* last file whose start <= offset */
  uint64_t SkipToRecordBegin(size_t file, uint64_t offset_in_file);
  void ReadRange(uint64_t offset, char *dst, size_t n);      /* across files, multi-threaded */
  void ReadRangeSerial(uint64_t offset, char *dst, size_t n);

  std::vector<File> files_;
  std::vector<uint64_t> start_;   /* start_[i] = stream offset of file i; start_[n] = T */
  uint64_t begin_, end_, cur_;
  int read_threads_;
  b200p::HostPool *pool_;   /* readers of ReadRange, created on first use */
  /* NextChunk / NextRecord state */
  std::vector<char> chunk_;
  size_t chunk_size_hint_;
  size_t rec_pos_, rec_end_;
};

/*
 * "stdin": the lines arriving on file descriptor 0, as one unpartitioned stream (the job of the
 * reference's SingleFileSplit, src/io/single_file_split.h:27-170, reached through
 * InputSplit::Create("stdin", ...), src/io.cc:96-98).  BeforeFirst rewinds when stdin is a
 * regular file and is a no-op on a pipe, where the reference's fseek fails just as silently.
 */
class StdinSplit : public InputSplit, public LineSource {
 public:
  StdinSplit();
  virtual void HintChunkSize(size_t chunk_size);
  virtual size_t GetTotalSize(void);
  virtual void BeforeFirst(void);
  virtual bool NextRecord(Blob *out_rec);
  virtual bool NextChunk(Blob *out_chunk);
  virtual void ResetPartition(unsigned part_index, unsigned num_parts);
  virtual size_t ReadLines(void *buf, size_t cap);
  virtual uint64_t PassBytes() const { return 0; }

 private:
  std::vector<char> carry_;   /* bytes read from the descriptor but not handed out yet */
  bool eof_;
  std::vector<char> chunk_;
  size_t chunk_size_hint_;
  size_t rec_pos_, rec_end_;
};

/*
 * "uri#cachefile" for InputSplit::Create: the first pass over `base` also writes every chunk to
 * the cache file; from the first BeforeFirst() on -- or from the start, when the file already
 * exists -- chunks come from the cache and the original files are not touched again (the job of
 * the reference's CachedInputSplit, src/io/cached_input_split.h:31-187).  The file is the
 * reference's: per chunk a native size_t byte count, then the bytes (:157-160, :176-186), so
 * caches can be swapped between the two implementations.
 */
class CachedSplit : public InputSplit {
 public:
  CachedSplit(InputSplit *base, const std::string &cache_file);
  virtual ~CachedSplit();
  virtual void HintChunkSize(size_t chunk_size) { base_->HintChunkSize(chunk_size); }
  virtual size_t GetTotalSize(void) { return base_->GetTotalSize(); }
  virtual void BeforeFirst(void);
  virtual bool NextRecord(Blob *out_rec);
  virtual bool NextChunk(Blob *out_chunk);
  virtual void ResetPartition(unsigned part_index, unsigned num_parts);

 private:
  bool OpenForReplay();
  InputSplit *base_;
  std::string cache_file_;
  std::FILE *writer_;   /* non-NULL while the first pass is being recorded */
  std::FILE *reader_;   /* non-NULL once chunks come from the cache */
  std::vector<char> chunk_;
  size_t rec_pos_, rec_end_;
};

}  // namespace b200
}  // namespace dmlc
#endif  // DMLC_CORE_B200_TEXT_SPLIT_H_
