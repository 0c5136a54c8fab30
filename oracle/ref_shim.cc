/*
 * TEST INFRASTRUCTURE ONLY -- never linked into the product library.
 *
 * A thin extern "C" harness around the UNMODIFIED reference (rahul003/dmlc-core).
 * It is compiled by oracle/Makefile together with the reference's own .cc files,
 * from where they lie under /root/reference, into oracle/_ref/libdmlc_ref.so.
 * No reference source is copied into this repository; this file only *calls* it:
 *
 *   refshim_parse_uri     dmlc::Parser<I>::Create + Next()/Value() drain
 *                         (reference src/data.cc:62-85,131-147; src/data/parser.h:95-109)
 *   refshim_parse_buffer  <Format>Parser<I>::ParseBlock on an in-memory block
 *                         (reference src/data/libsvm_parser.h:35-99, libfm_parser.h:35-93,
 *                          csv_parser.h:63-104)
 *   refshim_strtof/...    dmlc::data::strtof / strtouint (reference src/data/strtonum.h:37-150)
 *   refshim_split_bytes   InputSplit::Create(..., "text") + NextChunk drain
 *                         (reference src/io.cc:75-131, src/io/input_split_base.cc:30-64)
 *   refshim_time_uri      the drain loop of test/libsvm_parser_test.cc:14-34, timed
 *
 * Uses: validating oracle/dmlc_oracle.c, generating tests/golden/, and the
 * `bench.py --impl reference` / cpu_baseline leg.
 */
#include <dmlc/data.h>
#include <dmlc/io.h>
#include <random>   /* the reference header uses std::mt19937 without including it */
#include <dmlc/input_split_shuffle.h>
#include <dmlc/omp.h>
#include <dmlc/timer.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "data/csv_parser.h"
#include "data/libfm_parser.h"
#include "data/libsvm_parser.h"
#include "data/strtonum.h"
#include "parse_result.h"

namespace {

template <typename I>
struct Accum {
  std::vector<uint64_t> offset;
  std::vector<float> label, weight, value;
  std::vector<uint64_t> qid;
  std::vector<I> field, index;
  uint64_t nblocks;
  Accum() : nblocks(0) { offset.push_back(0); }
  void Add(const dmlc::RowBlock<I> &b) {
    ++nblocks;
    const size_t base = b.offset[0];
    const size_t nnz = b.offset[b.size] - base;
    const uint64_t run = index.size();
    for (size_t i = 1; i <= b.size; ++i) offset.push_back(run + (b.offset[i] - base));
    label.insert(label.end(), b.label, b.label + b.size);
    /* weight / qid may legally be shorter than `size` in the reference (no check);
       the containers hide their true length behind a raw pointer, so the shim can
       only take `size` entries when the pointer is non-NULL. Generators never mix. */
    if (b.weight != NULL) weight.insert(weight.end(), b.weight, b.weight + b.size);
    if (b.qid != NULL) qid.insert(qid.end(), b.qid, b.qid + b.size);
    if (b.field != NULL) field.insert(field.end(), b.field + base, b.field + base + nnz);
    index.insert(index.end(), b.index + base, b.index + base + nnz);
    if (b.value != NULL) value.insert(value.end(), b.value + base, b.value + base + nnz);
  }
};

template <typename T>
T *Dup(const std::vector<T> &v) {
  if (v.empty()) return NULL;
  T *p = static_cast<T *>(std::malloc(v.size() * sizeof(T)));
  std::memcpy(p, v.data(), v.size() * sizeof(T));
  return p;
}

template <typename I>
void Export(const Accum<I> &a, oracle_result *out) {
  out->n_offset = a.offset.size();
  out->n_label = a.label.size();
  out->n_weight = a.weight.size();
  out->n_qid = a.qid.size();
  out->n_field = a.field.size();
  out->n_index = a.index.size();
  out->n_value = a.value.size();
  out->offset = Dup(a.offset);
  out->label = Dup(a.label);
  out->weight = Dup(a.weight);
  out->qid = Dup(a.qid);
  out->field = Dup(a.field);
  out->index = Dup(a.index);
  out->value = Dup(a.value);
  out->index_bytes = sizeof(I);
  out->n_blocks = a.nblocks;
}

void SetError(oracle_result *out, const char *what) {
  out->error = 1;
  std::snprintf(out->errmsg, sizeof(out->errmsg), "%s", what);
}

template <typename I>
void ParseUri(const char *uri, unsigned part, unsigned nparts, const char *type,
              oracle_result *out) {
  Accum<I> acc;
  dmlc::Parser<I> *p = NULL;
  try {
    p = dmlc::Parser<I>::Create(uri, part, nparts, type);
    while (p->Next()) acc.Add(p->Value());
    out->bytes_read = p->BytesRead();
  } catch (const dmlc::Error &e) {
    SetError(out, e.what());
  } catch (const std::exception &e) {
    SetError(out, e.what());
  }
  delete p;
  Export(acc, out);
}

/* ParseBlock is protected in the reference parsers: open it up by derivation. */
template <typename I>
struct OpenLibSVM : public dmlc::data::LibSVMParser<I> {
  OpenLibSVM() : dmlc::data::LibSVMParser<I>(NULL, 1) {}
  void Run(char *b, char *e, dmlc::data::RowBlockContainer<I> *o) { this->ParseBlock(b, e, o); }
};
template <typename I>
struct OpenLibFM : public dmlc::data::LibFMParser<I> {
  OpenLibFM() : dmlc::data::LibFMParser<I>(NULL, 1) {}
  void Run(char *b, char *e, dmlc::data::RowBlockContainer<I> *o) { this->ParseBlock(b, e, o); }
};
template <typename I>
struct OpenCSV : public dmlc::data::CSVParser<I> {
  explicit OpenCSV(const std::map<std::string, std::string> &args)
      : dmlc::data::CSVParser<I>(NULL, args, 1) {}
  void Run(char *b, char *e, dmlc::data::RowBlockContainer<I> *o) { this->ParseBlock(b, e, o); }
};

template <typename I>
void ParseBuffer(int format, int label_column, const char *buf, size_t len, oracle_result *out) {
  /* Private copy with NUL padding: the reference converters ignore their `end`
     argument (strtonum.h:211-216) and may look a few bytes past the block. */
  std::vector<char> copy(len + 64, '\0');
  if (len != 0) std::memcpy(copy.data(), buf, len);
  Accum<I> acc;
  dmlc::data::RowBlockContainer<I> c;
  try {
    if (format == ORACLE_LIBSVM) {
      OpenLibSVM<I> p;
      p.Run(copy.data(), copy.data() + len, &c);
    } else if (format == ORACLE_LIBFM) {
      OpenLibFM<I> p;
      p.Run(copy.data(), copy.data() + len, &c);
    } else {
      std::map<std::string, std::string> args;
      char tmp[32];
      std::snprintf(tmp, sizeof(tmp), "%d", label_column);
      args["label_column"] = tmp;
      OpenCSV<I> p(args);
      p.Run(copy.data(), copy.data() + len, &c);
    }
    if (c.Size() != 0) acc.Add(c.GetBlock());
    /* keep the true weight/qid lengths, which GetBlock() hides */
    acc.weight = c.weight;
    acc.qid = c.qid;
  } catch (const dmlc::Error &e) {
    SetError(out, e.what());
  }
  Export(acc, out);
  out->bytes_read = len;
}

/* RowBlockIter<I>::Create(uri, ...) drained once (uri may carry "#cachefile": the reference's
   DiskRowIter, src/data/disk_row_iter.h). *num_col receives NumCol(). */
template <typename I>
void IterUri(const char *uri, unsigned part, unsigned nparts, const char *type,
                    oracle_result *out, uint64_t *num_col) {
  Accum<I> acc;
  dmlc::RowBlockIter<I> *it = NULL;
  try {
    it = dmlc::RowBlockIter<I>::Create(uri, part, nparts, type);
    it->BeforeFirst();
    while (it->Next()) acc.Add(it->Value());
    *num_col = it->NumCol();
  } catch (const std::exception &e) {
    SetError(out, e.what());
  }
  delete it;
  Export(acc, out);
}

}  // namespace

static inline uint64_t ShimFnv(uint64_t h, const void *data, size_t nbytes) {
  const unsigned char *p = static_cast<const unsigned char *>(data);
  for (size_t i = 0; i < nbytes; ++i) h = (h ^ p[i]) * 1099511628211ull;
  return h;
}
template <typename I>
static void ShimHash(dmlc::Parser<I> *p, uint64_t *hash7, uint64_t *rows, uint64_t *nnz) {
  for (int k = 0; k < 7; ++k) hash7[k] = 14695981039346656037ull;
  while (p->Next()) {
    const dmlc::RowBlock<I> &b = p->Value();
    const size_t base = b.offset[0], n = b.offset[b.size] - base;
    for (size_t i = 0; i < b.size; ++i) {
      const uint64_t len = b.offset[i + 1] - b.offset[i];
      hash7[0] = ShimFnv(hash7[0], &len, 8);
    }
    hash7[1] = ShimFnv(hash7[1], b.label, b.size * sizeof(float));
    if (b.weight != NULL) hash7[2] = ShimFnv(hash7[2], b.weight, b.size * sizeof(float));
    if (b.qid != NULL) hash7[3] = ShimFnv(hash7[3], b.qid, b.size * sizeof(uint64_t));
    if (b.field != NULL) hash7[4] = ShimFnv(hash7[4], b.field + base, n * sizeof(I));
    if (b.index != NULL) hash7[5] = ShimFnv(hash7[5], b.index + base, n * sizeof(I));
    if (b.value != NULL) hash7[6] = ShimFnv(hash7[6], b.value + base, n * sizeof(float));
    *rows += b.size;
    *nnz += n;
  }
}

extern "C" {

void refshim_parse_uri(const char *uri, unsigned part, unsigned nparts, const char *type,
                       int index_bytes, oracle_result *out) {
  std::memset(out, 0, sizeof(*out));
  if (index_bytes == 8) {
    ParseUri<uint64_t>(uri, part, nparts, type, out);
  } else {
    ParseUri<uint32_t>(uri, part, nparts, type, out);
  }
}

void refshim_parse_buffer(int format, int index_bytes, int label_column, const char *buf,
                          size_t len, oracle_result *out) {
  std::memset(out, 0, sizeof(*out));
  if (index_bytes == 8) {
    ParseBuffer<uint64_t>(format, label_column, buf, len, out);
  } else {
    ParseBuffer<uint32_t>(format, label_column, buf, len, out);
  }
}

void refshim_iter_uri(const char *uri, unsigned part, unsigned nparts, const char *type,
                      int index_bytes, oracle_result *out, uint64_t *num_col) {
  std::memset(out, 0, sizeof(*out));
  *num_col = 0;
  if (index_bytes == 8) {
    IterUri<uint64_t>(uri, part, nparts, type, out, num_col);
  } else {
    IterUri<uint32_t>(uri, part, nparts, type, out, num_col);
  }
}

void refshim_free(oracle_result *r) {
  std::free(r->offset);
  std::free(r->label);
  std::free(r->weight);
  std::free(r->qid);
  std::free(r->field);
  std::free(r->index);
  std::free(r->value);
  std::memset(r, 0, sizeof(*r));
}

/* s must be NUL terminated. Returns the IEEE bits; *consumed = endptr - s. */
uint32_t refshim_strtof_bits(const char *s, int64_t *consumed) {
  char *end = NULL;
  float v = dmlc::data::strtof(s, &end);
  if (consumed != NULL) *consumed = end - s;
  uint32_t bits;
  std::memcpy(&bits, &v, sizeof(bits));
  return bits;
}

/* width 4: Str2T<uint32_t> path (through int, strtonum.h:186-190); width 8: uint64_t.
   *threw = 1 when the reference raised dmlc::Error (negative sign). */
uint64_t refshim_strtouint(const char *s, int width, int *threw) {
  *threw = 0;
  try {
    if (width == 8) return dmlc::data::Str2T<uint64_t>::get(s, s + std::strlen(s));
    return dmlc::data::Str2T<uint32_t>::get(s, s + std::strlen(s));
  } catch (const dmlc::Error &) {
    *threw = 1;
    return 0;
  }
}

/* Concatenated bytes of every chunk InputSplit "text" yields for (part, nparts).
   Returns the total length; copies at most `cap` bytes into buf. */
uint64_t refshim_split_bytes(const char *uri, unsigned part, unsigned nparts, char *buf,
                             uint64_t cap, uint64_t *nchunks) {
  uint64_t total = 0, n = 0;
  dmlc::InputSplit *split = dmlc::InputSplit::Create(uri, part, nparts, "text");
  dmlc::InputSplit::Blob blob;
  while (split->NextChunk(&blob)) {
    if (total < cap) {
      uint64_t take = blob.size;
      if (total + take > cap) take = cap - total;
      std::memcpy(buf + total, blob.dptr, take);
    }
    total += blob.size;
    ++n;
  }
  delete split;
  if (nchunks != NULL) *nchunks = n;
  return total;
}

/* FNV-1a of the seven arrays of Parser<I>::Create(uri, ...) drained once (row lengths instead of
   offsets, so the hash does not depend on how rows are cut into blocks): bench.py compares it
   with the same hash taken through the B200 parser on the same file. Returns 0, or -1 on error. */
/* reference InputSplitShuffle (include/dmlc/input_split_shuffle.h) drained with NextChunk over
   `passes` passes: the bytes in the order they come out */
uint64_t refshim_shuffle_bytes(const char *uri, unsigned part, unsigned nparts, unsigned shuffle_parts,
                               int seed, int passes, char *buf, uint64_t cap) {
  uint64_t total = 0;
  try {
    dmlc::InputSplit *split = dmlc::InputSplitShuffle::Create(uri, part, nparts, "text", shuffle_parts, seed);
    dmlc::InputSplit::Blob blob;
    for (int k = 0; k < passes; ++k) {
      if (k != 0) split->BeforeFirst();
      while (split->NextChunk(&blob)) {
        if (total + blob.size <= cap) std::memcpy(buf + total, blob.dptr, blob.size);
        total += blob.size;
      }
    }
    delete split;
  } catch (const std::exception &e) {
    std::fprintf(stderr, "refshim_shuffle_bytes: %s\n", e.what());
    return ~0ull;
  }
  return total;
}

int refshim_hash_uri(const char *uri, unsigned part, unsigned nparts, const char *type,
                     int index_bytes, int nthread, uint64_t *hash7, uint64_t *rows, uint64_t *nnz) {
  if (nthread > 0) omp_set_num_threads(nthread);
  *rows = 0;
  *nnz = 0;
  try {
    if (index_bytes == 8) {
      dmlc::Parser<uint64_t> *p = dmlc::Parser<uint64_t>::Create(uri, part, nparts, type);
      ShimHash(p, hash7, rows, nnz);
      delete p;
    } else {
      dmlc::Parser<uint32_t> *p = dmlc::Parser<uint32_t>::Create(uri, part, nparts, type);
      ShimHash(p, hash7, rows, nnz);
      delete p;
    }
  } catch (const std::exception &e) {
    std::fprintf(stderr, "refshim_hash_uri: %s\n", e.what());
    return -1;
  }
  return 0;
}

/* Timed drain, touching only sizes (test/libsvm_parser_test.cc:21-33 semantics).
   nthread > 0 sets the OpenMP team size FillData picks up (text_parser.h:98).
   Returns seconds for `passes` passes (BeforeFirst between them); -1 on error. */
double refshim_time_uri(const char *uri, unsigned part, unsigned nparts, const char *type,
                        int index_bytes, int nthread, int passes, uint64_t *rows,
                        uint64_t *nnz, uint64_t *bytes) {
  if (nthread > 0) omp_set_num_threads(nthread);
  uint64_t r = 0, z = 0, b = 0;
  double t0 = 0, t1 = 0;
  try {
    if (index_bytes == 8) {
      dmlc::Parser<uint64_t> *p = dmlc::Parser<uint64_t>::Create(uri, part, nparts, type);
      t0 = dmlc::GetTime();
      for (int k = 0; k < passes; ++k) {
        if (k != 0) p->BeforeFirst();
        while (p->Next()) {
          const dmlc::RowBlock<uint64_t> &v = p->Value();
          r += v.size;
          z += v.offset[v.size] - v.offset[0];
        }
      }
      t1 = dmlc::GetTime();
      b = p->BytesRead();
      delete p;
    } else {
      dmlc::Parser<uint32_t> *p = dmlc::Parser<uint32_t>::Create(uri, part, nparts, type);
      t0 = dmlc::GetTime();
      for (int k = 0; k < passes; ++k) {
        if (k != 0) p->BeforeFirst();
        while (p->Next()) {
          const dmlc::RowBlock<uint32_t> &v = p->Value();
          r += v.size;
          z += v.offset[v.size] - v.offset[0];
        }
      }
      t1 = dmlc::GetTime();
      b = p->BytesRead();
      delete p;
    }
  } catch (const std::exception &e) {
    std::fprintf(stderr, "refshim_time_uri: %s\n", e.what());
    return -1.0;
  }
  if (rows != NULL) *rows = r;
  if (nnz != NULL) *nnz = z;
  if (bytes != NULL) *bytes = b;
  return t1 - t0;
}

int refshim_max_threads(void) { return omp_get_max_threads(); }

}  // extern "C"
