#include "./text_split.h"

#include <dirent.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <cerrno>
#include <cstdlib>
#include <cstring>
#include <regex>
#include <thread>

#if defined(__SSE2__)
#include <emmintrin.h>
#endif

#include "../csrc/host_pool.h"
#include "./uri_spec.h"

namespace dmlc {
namespace b200 {

namespace {
inline bool IsEol(char c) { return c == '\n' || c == '\r'; }

std::string StripProtocol(const std::string &p) {
  if (p.compare(0, 7, "file://") == 0) return p.substr(7);
  CHECK(p.find("://") == std::string::npos)
      << "dmlc_core_b200 reads local files only, got " << p;
  return p;
}
std::string StripTrailingSlash(std::string s) {
  while (s.size() > 1 && s[s.size() - 1] == '/') s.resize(s.size() - 1);
  return s;
}
}  // namespace

TextSplit::TextSplit(const char *uri, unsigned part_index, unsigned num_parts)
    : begin_(0), end_(0), cur_(0), read_threads_(1), pool_(NULL), chunk_size_hint_(8u << 20),
      rec_pos_(0), rec_end_(0) {
  /* three quarters of the cores, at most 12 (measured on a 16-core B200 host: one 64 MiB chunk
     from the page cache at 39 GB/s with 8 readers, 48 GB/s with 12), shared fairly when several
     ranks run on one host (torchrun exports LOCAL_WORLD_SIZE) */
  unsigned hw = std::thread::hardware_concurrency();
  const char *local = std::getenv("LOCAL_WORLD_SIZE");
  const unsigned ranks = (local != NULL && std::atoi(local) > 0) ? static_cast<unsigned>(std::atoi(local)) : 1u;
  read_threads_ = static_cast<int>(std::max(1u, std::min(12u, hw * 3u / 4u / ranks)));
  const char *forced = std::getenv("DMLC_B200_READ_THREADS");
  if (forced != NULL && std::atoi(forced) > 0) read_threads_ = std::min(std::atoi(forced), 64);
  ExpandUri(uri);
  CHECK_NE(files_.size(), 0U) << "Cannot find any files that matches the URI pattern " << uri;
  start_.resize(files_.size() + 1);
  start_[0] = 0;
  for (size_t i = 0; i < files_.size(); ++i) start_[i + 1] = start_[i] + files_[i].size;
  ResetPartition(part_index, num_parts);
}

TextSplit::~TextSplit() {
  delete pool_;
  for (size_t i = 0; i < files_.size(); ++i) {
    if (files_[i].map != NULL) munmap(const_cast<char *>(files_[i].map), static_cast<size_t>(files_[i].size));
    if (files_[i].fd >= 0) close(files_[i].fd);
  }
}

/* One entry of the ';' list: a file, a directory (its non-empty regular files, listing order), or
   a regular expression matched against the full paths inside its directory
   (reference src/io/input_split_base.cc:96-175). */
void TextSplit::ExpandUri(const std::string &uri) {
  size_t pos = 0;
  while (pos <= uri.size()) {
    size_t semi = uri.find(';', pos);
    if (semi == std::string::npos) semi = uri.size();
    if (semi > pos) AddPath(StripProtocol(uri.substr(pos, semi - pos)), false);
    pos = semi + 1;
  }
}

void TextSplit::AddPath(const std::string &path, bool from_listing) {
  struct stat sb;
  if (stat(path.c_str(), &sb) == 0) {
    if (S_ISDIR(sb.st_mode)) {
      if (from_listing) return;  /* not recursive */
      DIR *dir = opendir(path.c_str());
      CHECK(dir != NULL) << "ListDirectory " << path << " error: " << strerror(errno);
      struct dirent *ent;
      while ((ent = readdir(dir)) != NULL) {
        if (!strcmp(ent->d_name, ".") || !strcmp(ent->d_name, "..")) continue;
        std::string child = path;
        if (child[child.size() - 1] != '/') child += '/';
        AddPath(child + ent->d_name, true);
      }
      closedir(dir);
      return;
    }
    if (sb.st_size == 0) return;  /* empty files are dropped */
    File f;
    f.path = path;
    f.size = static_cast<uint64_t>(sb.st_size);
    f.fd = open(path.c_str(), O_RDONLY);
    CHECK(f.fd >= 0) << "cannot open " << path << ": " << strerror(errno);
    f.map = NULL;
    /* The readers copy out of a read-only mapping of the file with streaming stores instead of
       calling pread: no read-for-ownership of the staging buffer's lines, no system call per
       piece (Parser::Create from the page cache: 36-38 GB/s instead of 30-31 on the 16-core
       bench host).  DMLC_B200_READ_MMAP=0 keeps pread -- a file that is
       truncated while it is being read then yields an error instead of a SIGBUS. */
    const char *want_map = std::getenv("DMLC_B200_READ_MMAP");
    if (want_map == NULL || want_map[0] != '0') {
      void *m = mmap(NULL, static_cast<size_t>(f.size), PROT_READ, MAP_SHARED, f.fd, 0);
      if (m != MAP_FAILED) {
        madvise(m, static_cast<size_t>(f.size), MADV_SEQUENTIAL);
        f.map = static_cast<const char *>(m);
      }
    }
    files_.push_back(f);
    return;
  }
  if (from_listing) return;
  /* not an existing path: treat the last component as a regex over its directory */
  const size_t slash = path.rfind('/');
  CHECK(slash != std::string::npos && slash + 1 != path.size())
      << "LocalFileSystem.GetPathInfo: " << path << " error: " << strerror(ENOENT);
  const std::string dirname = slash == 0 ? "/" : path.substr(0, slash);
  DIR *dir = opendir(dirname.c_str());
  CHECK(dir != NULL) << "ListDirectory " << dirname << " error: " << strerror(errno);
  std::regex pattern;
  try {
    pattern = std::regex(path);
  } catch (const std::regex_error &e) {
    closedir(dir);
    LOG(FATAL) << e.what() << " bad regex " << path;
  }
  struct dirent *ent;
  while ((ent = readdir(dir)) != NULL) {
    if (!strcmp(ent->d_name, ".") || !strcmp(ent->d_name, "..")) continue;
    std::string child = dirname;
    if (child[child.size() - 1] != '/') child += '/';
    child += ent->d_name;
    if (std::regex_match(StripTrailingSlash(child), pattern)) AddPath(child, true);
  }
  closedir(dir);
}

size_t TextSplit::GetTotalSize(void) { return static_cast<size_t>(start_.back()); }

size_t TextSplit::FileAt(uint64_t offset) const {
  /* upper_bound - 1 over the prefix sums, as src/io/input_split_base.cc:38-43 */
  return static_cast<size_t>(std::upper_bound(start_.begin(), start_.end(), offset) -
                             start_.begin()) - 1;
}

/* Bytes to skip from (file, offset) to the next record begin: through the first EOL and every
   EOL byte directly after it, stopping at the end of that file (line_split.cc:9-26). */
uint64_t TextSplit::SkipToRecordBegin(size_t file, uint64_t off) {
  const File &f = files_[file];
  char buf[4096];
  uint64_t n = 0;
  bool in_eol_run = false;
  for (;;) {
    const ssize_t got = pread(f.fd, buf, sizeof(buf), static_cast<off_t>(off + n));
    if (got <= 0) return n;
    for (ssize_t i = 0; i < got; ++i) {
      if (!in_eol_run) {
        ++n;
        if (IsEol(buf[i])) in_eol_run = true;
      } else {
        if (!IsEol(buf[i])) return n;
        ++n;
      }
    }
  }
}

void TextSplit::ResetPartition(unsigned rank, unsigned nsplit) {
  CHECK(rank < nsplit) << "invalid input parameter for InputSplit::Create";
  const uint64_t total = start_.back();
  const uint64_t step = (total + nsplit - 1) / nsplit;
  begin_ = std::min<uint64_t>(step * rank, total);
  end_ = std::min<uint64_t>(step * (rank + 1), total);
  if (begin_ != end_) {
    const size_t fe = FileAt(end_);
    if (end_ != start_[fe]) end_ += SkipToRecordBegin(fe, end_ - start_[fe]);
    const size_t fb = FileAt(begin_);
    if (begin_ != start_[fb]) begin_ += SkipToRecordBegin(fb, begin_ - start_[fb]);
  }
  BeforeFirst();
}

void TextSplit::BeforeFirst(void) {
  cur_ = begin_;
  rec_pos_ = rec_end_ = 0;
}

void TextSplit::HintChunkSize(size_t chunk_size) {
  chunk_size_hint_ = std::max(chunk_size_hint_, chunk_size);
}

/* copy with stores that bypass the cache: the destination is a staging buffer the CPU never reads */
static void StreamCopy(char *dst, const char *src, size_t n) {
#if defined(__SSE2__)
  size_t head = (16 - (reinterpret_cast<uintptr_t>(dst) & 15)) & 15;
  if (head > n) head = n;
  std::memcpy(dst, src, head);
  dst += head; src += head; n -= head;
  size_t i = 0;
  for (; i + 64 <= n; i += 64) {
    const __m128i a = _mm_loadu_si128(reinterpret_cast<const __m128i *>(src + i));
    const __m128i b = _mm_loadu_si128(reinterpret_cast<const __m128i *>(src + i + 16));
    const __m128i c = _mm_loadu_si128(reinterpret_cast<const __m128i *>(src + i + 32));
    const __m128i d = _mm_loadu_si128(reinterpret_cast<const __m128i *>(src + i + 48));
    _mm_stream_si128(reinterpret_cast<__m128i *>(dst + i), a);
    _mm_stream_si128(reinterpret_cast<__m128i *>(dst + i + 16), b);
    _mm_stream_si128(reinterpret_cast<__m128i *>(dst + i + 32), c);
    _mm_stream_si128(reinterpret_cast<__m128i *>(dst + i + 48), d);
  }
  _mm_sfence();
  std::memcpy(dst + i, src + i, n - i);
#else
  std::memcpy(dst, src, n);
#endif
}

void TextSplit::ReadRangeSerial(uint64_t offset, char *dst, size_t n) {
  size_t f = FileAt(offset);
  while (n != 0) {
    CHECK(f < files_.size()) << "read past the end of the input files";
    const uint64_t in_file = offset - start_[f];
    const size_t want = static_cast<size_t>(std::min<uint64_t>(n, files_[f].size - in_file));
    if (files_[f].map != NULL) {
      StreamCopy(dst, files_[f].map + in_file, want);
      dst += want;
      offset += want;
      n -= want;
      ++f;
      continue;
    }
    size_t done = 0;
    while (done < want) {
      const ssize_t got = pread(files_[f].fd, dst + done, want - done,
                                static_cast<off_t>(in_file + done));
      CHECK(got > 0) << "short read on " << files_[f].path << ": "
                     << (got < 0 ? strerror(errno) : "file shrank");
      done += static_cast<size_t>(got);
    }
    dst += want;
    offset += want;
    n -= want;
    ++f;
  }
}

void TextSplit::ReadRange(uint64_t offset, char *dst, size_t n) {
  const size_t kMinPiece = 2u << 20;
  const int pieces = static_cast<int>(std::min<size_t>(4 * static_cast<size_t>(read_threads_),
                                                       (n + kMinPiece - 1) / kMinPiece));
  if (pieces <= 1 || read_threads_ <= 1) {
    ReadRangeSerial(offset, dst, n);
    return;
  }
  /* persistent readers: spawning a dozen threads per 64 MiB chunk costs a fifth of the read */
  if (pool_ == NULL || pool_->threads() != read_threads_) {
    delete pool_;
    pool_ = new b200p::HostPool(read_threads_);
  }
  const size_t per = ((n + pieces - 1) / pieces + 4095) / 4096 * 4096;
  std::vector<std::string> errors(pieces);
  pool_->Run(pieces, [&](int i) {
    const size_t lo = std::min(n, per * i), hi = std::min(n, per * (i + 1));
    if (lo == hi) return;
    try {
      ReadRangeSerial(offset + lo, dst + lo, hi - lo);
    } catch (const dmlc::Error &e) {
      errors[i] = e.what();
    }
  });
  for (int i = 0; i < pieces; ++i) {
    if (!errors[i].empty()) LOG(FATAL) << errors[i];
  }
}

size_t TextSplit::ReadLines(void *buf, size_t cap) {
  if (cur_ >= end_ || cap == 0) return 0;
  char *dst = static_cast<char *>(buf);
  const uint64_t left = end_ - cur_;
  if (left <= cap) {
    ReadRange(cur_, dst, static_cast<size_t>(left));
    cur_ = end_;
    return static_cast<size_t>(left);
  }
  ReadRange(cur_, dst, cap);
  /* cut after the last EOL byte (FindLastRecordBegin, line_split.cc:27-34); the bytes behind
     it are simply read again by the next call */
  size_t cut = cap;
  while (cut > 1 && !IsEol(dst[cut - 1])) --cut;
  if (cut <= 1 && !IsEol(dst[0])) return kLineTooLong;
  if (cut <= 1) cut = 1;
  cur_ += cut;
  return cut;
}

bool TextSplit::NextChunk(Blob *out_chunk) {
  size_t cap = chunk_size_hint_;
  for (;;) {
    if (chunk_.size() < cap + 1) chunk_.resize(cap + 1);
    const size_t n = ReadLines(chunk_.data(), cap);
    if (n == kLineTooLong) {
      cap *= 2;  /* one line longer than the buffer: grow, like Chunk::Load (:249-251) */
      continue;
    }
    if (n == 0) return false;
    chunk_[n] = '\0';
    out_chunk->dptr = chunk_.data();
    out_chunk->size = n;
    rec_pos_ = rec_end_ = 0;
    return true;
  }
}

/* records are served out of chunks (reference LineSplitter::ExtractNextRecord,
   src/io/line_split.cc:36-55): a record is the line plus the EOL run that follows it; the last
   byte of that run (or the byte after the chunk) is overwritten by NUL.  `base` has one spare
   byte behind `end`. */
static void CutRecord(char *base, size_t *pos, size_t end, InputSplit::Blob *out_rec) {
  size_t p = *pos;
  while (p != end && !IsEol(base[p])) ++p;
  while (p != end && IsEol(base[p])) ++p;
  if (p == end) base[p] = '\0'; else base[p - 1] = '\0';
  out_rec->dptr = base + *pos;
  out_rec->size = p - *pos;
  *pos = p;
}

bool TextSplit::NextRecord(Blob *out_rec) {
  while (rec_pos_ == rec_end_) {
    Blob chunk;
    if (!NextChunk(&chunk)) return false;
    rec_pos_ = 0;
    rec_end_ = chunk.size;
  }
  CutRecord(chunk_.data(), &rec_pos_, rec_end_, out_rec);
  return true;
}

/* ------------------------------------------------------------------------- stdin ---------- */
StdinSplit::StdinSplit() : eof_(false), chunk_size_hint_(8u << 20), rec_pos_(0), rec_end_(0) {}

void StdinSplit::HintChunkSize(size_t chunk_size) {
  chunk_size_hint_ = std::max<size_t>(chunk_size, 1u << 18);
}

size_t StdinSplit::GetTotalSize(void) {
  struct stat st;
  return fstat(0, &st) == 0 ? static_cast<size_t>(st.st_size) : 0;
}

void StdinSplit::BeforeFirst(void) {
  if (lseek(0, 0, SEEK_SET) == 0) {   /* a redirected file; a pipe cannot be replayed */
    carry_.clear();
    eof_ = false;
  }
  rec_pos_ = rec_end_ = 0;
}

void StdinSplit::ResetPartition(unsigned part_index, unsigned num_parts) {
  CHECK(part_index == 0 && num_parts == 1) << "stdin cannot be partitioned";
  BeforeFirst();
}

size_t StdinSplit::ReadLines(void *buf, size_t cap) {
  char *dst = static_cast<char *>(buf);
  if (carry_.size() >= cap && cap != 0) {
    /* more is waiting than fits: serve whole lines out of the carry alone */
    size_t cut = cap;
    while (cut > 0 && !IsEol(carry_[cut - 1])) --cut;
    if (cut == 0) return kLineTooLong;
    std::memcpy(dst, carry_.data(), cut);
    carry_.erase(carry_.begin(), carry_.begin() + cut);
    return cut;
  }
  size_t have = carry_.size();
  if (have != 0) std::memcpy(dst, carry_.data(), have);
  carry_.clear();
  while (!eof_ && have < cap) {
    const ssize_t got = read(0, dst + have, cap - have);
    if (got < 0 && errno == EINTR) continue;
    CHECK(got >= 0) << "read from stdin failed: " << strerror(errno);
    if (got == 0) eof_ = true;
    have += static_cast<size_t>(got);
  }
  if (eof_) return have;            /* the rest of the stream, with or without a final EOL */
  size_t cut = have;
  while (cut > 0 && !IsEol(dst[cut - 1])) --cut;
  carry_.assign(dst + cut, dst + have);   /* the unfinished line waits for the next call */
  if (cut == 0) return kLineTooLong;
  return cut;
}

bool StdinSplit::NextChunk(Blob *out_chunk) {
  size_t cap = chunk_size_hint_;
  for (;;) {
    if (chunk_.size() < cap + 1) chunk_.resize(cap + 1);
    const size_t n = ReadLines(chunk_.data(), cap);
    if (n == kLineTooLong) {
      cap *= 2;
      continue;
    }
    if (n == 0) return false;
    chunk_[n] = '\0';
    out_chunk->dptr = chunk_.data();
    out_chunk->size = n;
    rec_pos_ = rec_end_ = 0;
    return true;
  }
}

bool StdinSplit::NextRecord(Blob *out_rec) {
  while (rec_pos_ == rec_end_) {
    Blob chunk;
    if (!NextChunk(&chunk)) return false;
    rec_pos_ = 0;
    rec_end_ = chunk.size;
  }
  CutRecord(chunk_.data(), &rec_pos_, rec_end_, out_rec);
  return true;
}

/* ------------------------------------------------------------------- "#cachefile" -------- */
CachedSplit::CachedSplit(InputSplit *base, const std::string &cache_file)
    : base_(base), cache_file_(cache_file), writer_(NULL), reader_(NULL), rec_pos_(0), rec_end_(0) {
  if (!OpenForReplay()) {
    writer_ = std::fopen(cache_file_.c_str(), "wb");
    CHECK(writer_ != NULL) << "cannot create cache file " << cache_file_;
  }
}

CachedSplit::~CachedSplit() {
  if (writer_ != NULL) std::fclose(writer_);
  if (reader_ != NULL) std::fclose(reader_);
  delete base_;
}

bool CachedSplit::OpenForReplay() {
  reader_ = std::fopen(cache_file_.c_str(), "rb");
  return reader_ != NULL;
}

void CachedSplit::ResetPartition(unsigned, unsigned) {
  LOG(FATAL) << "ResetPartition is not supported in CachedInputSplit";
}

void CachedSplit::BeforeFirst(void) {
  if (writer_ != NULL) {
    /* the recording pass must reach the end of the partition before the cache is complete */
    Blob rest;
    while (NextChunk(&rest)) {}
    std::fclose(writer_);
    writer_ = NULL;
    CHECK(OpenForReplay()) << "Failed to initialize CachedIter";
  } else {
    std::rewind(reader_);
  }
  rec_pos_ = rec_end_ = 0;
}

bool CachedSplit::NextChunk(Blob *out_chunk) {
  rec_pos_ = rec_end_ = 0;
  if (writer_ != NULL) {
    Blob blob;
    if (!base_->NextChunk(&blob)) return false;
    const size_t size = blob.size;
    CHECK(std::fwrite(&size, sizeof(size), 1, writer_) == 1 &&
          std::fwrite(blob.dptr, 1, size, writer_) == size) << "cannot write " << cache_file_;
    /* own copy with a spare byte, so that records can be cut out of it */
    chunk_.assign(static_cast<const char *>(blob.dptr), static_cast<const char *>(blob.dptr) + size);
    chunk_.push_back('\0');
    out_chunk->dptr = chunk_.data();
    out_chunk->size = size;
    return true;
  }
  size_t size = 0;
  const size_t got = std::fread(&size, 1, sizeof(size), reader_);
  if (got == 0) return false;
  CHECK(got == sizeof(size)) << cache_file_ << " has invalid cache file format";
  chunk_.resize(size + 1);
  CHECK(std::fread(chunk_.data(), 1, size, reader_) == size)
      << cache_file_ << " has invalid cache file format";
  chunk_[size] = '\0';
  out_chunk->dptr = chunk_.data();
  out_chunk->size = size;
  return true;
}

bool CachedSplit::NextRecord(Blob *out_rec) {
  while (rec_pos_ == rec_end_) {
    Blob chunk;
    if (!NextChunk(&chunk)) return false;
    rec_pos_ = 0;
    rec_end_ = chunk.size;
  }
  CutRecord(chunk_.data(), &rec_pos_, rec_end_, out_rec);
  return true;
}

}  // namespace b200

InputSplit *InputSplit::Create(const char *uri, unsigned part_index, unsigned num_parts,
                               const char *type) {
  b200::UriSpec spec(uri, part_index, num_parts);
  if (spec.uri == "stdin") return new b200::StdinSplit();   /* before the part check, like io.cc:96-99 */
  CHECK(part_index < num_parts) << "invalid input parameter for InputSplit::Create";
  CHECK(!strcmp(type, "text")) << "unknown input split type " << type
                               << " (dmlc_core_b200 provides the \"text\" split only)";
  InputSplit *files = new b200::TextSplit(spec.uri.c_str(), part_index, num_parts);
  if (spec.cache_file.empty()) return files;
  return new b200::CachedSplit(files, spec.cache_file);   /* io.cc:123-127 */
}

}  // namespace dmlc
