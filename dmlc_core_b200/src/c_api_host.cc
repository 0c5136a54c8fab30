/*
 * extern "C" access to the C++ host layer (dmlc::Parser / RowBlockIter / InputSplit as provided
 * by this package) for language bindings and for the Python test-suite, which drives the C++
 * API exactly like the reference's drivers do (reference test/libsvm_parser_test.cc:14-34,
 * test/csv_parser_test.cc:22-57, test/dataiter_test.cc).
 */
#include <dmlc/data.h>
#include <dmlc/io.h>
#include <random>   /* the reference's header (drop-in build) uses std::mt19937 without including it */
#include <dmlc/input_split_shuffle.h>
#include <dmlc/timer.h>

#include <cstdio>
#include <cstring>

extern "C" {

typedef struct dmlcb_block {
  uint64_t size;
  const uint64_t *offset;
  const float *label;
  const float *weight;
  const uint64_t *qid;
  const void *field;
  const void *index;
  const float *value;
} dmlcb_block;

struct dmlcb_parser {
  int index_bytes;
  dmlc::Parser<uint32_t> *p32;
  dmlc::Parser<uint64_t> *p64;
};

}  // extern "C"

template <typename I>
static void Export(const dmlc::RowBlock<I> &b, dmlcb_block *out) {
  out->size = b.size;
  out->offset = reinterpret_cast<const uint64_t *>(b.offset);
  out->label = b.label;
  out->weight = b.weight;
  out->qid = b.qid;
  out->field = b.field;
  out->index = b.index;
  out->value = b.value;
}

/* FNV-1a over the bytes of one array, carried across blocks */
static inline uint64_t Fnv(uint64_t h, const void *data, size_t nbytes) {
  const unsigned char *p = static_cast<const unsigned char *>(data);
  for (size_t i = 0; i < nbytes; ++i) h = (h ^ p[i]) * 1099511628211ull;
  return h;
}

template <typename I>
static void DrainTimed(dmlc::Parser<I> *p, int passes, int hash_first, uint64_t *rows, uint64_t *nnz,
                       uint64_t *hash7, double *seconds) {
  uint64_t r = 0, z = 0;
  if (hash_first) {
    /* one untimed pass hashing label, weight, qid, field, index, value and the row lengths:
       independent of how the rows are cut into blocks */
    for (int k = 0; k < 7; ++k) hash7[k] = 14695981039346656037ull;
    while (p->Next()) {
      const dmlc::RowBlock<I> &b = p->Value();
      const size_t base = b.offset[0], n = b.offset[b.size] - base;
      for (size_t i = 0; i < b.size; ++i) {
        const uint64_t len = b.offset[i + 1] - b.offset[i];
        hash7[0] = Fnv(hash7[0], &len, 8);
      }
      hash7[1] = Fnv(hash7[1], b.label, b.size * sizeof(float));
      if (b.weight != NULL) hash7[2] = Fnv(hash7[2], b.weight, b.size * sizeof(float));
      if (b.qid != NULL) hash7[3] = Fnv(hash7[3], b.qid, b.size * sizeof(uint64_t));
      if (b.field != NULL) hash7[4] = Fnv(hash7[4], b.field + base, n * sizeof(I));
      if (b.index != NULL) hash7[5] = Fnv(hash7[5], b.index + base, n * sizeof(I));
      if (b.value != NULL) hash7[6] = Fnv(hash7[6], b.value + base, n * sizeof(float));
    }
    p->BeforeFirst();
  }
  const double t0 = dmlc::GetTime();
  for (int k = 0; k < passes; ++k) {
    if (k != 0) p->BeforeFirst();
    while (p->Next()) {
      const dmlc::RowBlock<I> &v = p->Value();
      r += v.size;
      z += v.offset[v.size] - v.offset[0];
    }
  }
  *seconds = dmlc::GetTime() - t0;
  *rows = r;
  *nnz = z;
}

extern "C" {

static void Say(char *err, uint64_t errlen, const char *what) {
  if (err != NULL && errlen != 0) std::snprintf(err, errlen, "%s", what);
}

void *dmlcb_parser_create(const char *uri, unsigned part, unsigned nparts, const char *type,
                          int index_bytes, char *err, uint64_t errlen) {
  dmlcb_parser *h = new dmlcb_parser();
  h->index_bytes = index_bytes;
  h->p32 = NULL;
  h->p64 = NULL;
  try {
    if (index_bytes == 8) {
      h->p64 = dmlc::Parser<uint64_t>::Create(uri, part, nparts, type);
    } else {
      h->p32 = dmlc::Parser<uint32_t>::Create(uri, part, nparts, type);
    }
  } catch (const std::exception &e) {
    Say(err, errlen, e.what());
    delete h;
    return NULL;
  }
  return h;
}

/* 1: a block was produced; 0: end of data; -1: dmlc::Error (message in err) */
int dmlcb_parser_next(void *handle, dmlcb_block *out, char *err, uint64_t errlen) {
  dmlcb_parser *h = static_cast<dmlcb_parser *>(handle);
  try {
    if (h->p64 != NULL) {
      if (!h->p64->Next()) return 0;
      Export(h->p64->Value(), out);
    } else {
      if (!h->p32->Next()) return 0;
      Export(h->p32->Value(), out);
    }
  } catch (const std::exception &e) {
    Say(err, errlen, e.what());
    return -1;
  }
  return 1;
}

void dmlcb_parser_before_first(void *handle) {
  dmlcb_parser *h = static_cast<dmlcb_parser *>(handle);
  if (h->p64 != NULL) h->p64->BeforeFirst(); else h->p32->BeforeFirst();
}

uint64_t dmlcb_parser_bytes_read(void *handle) {
  dmlcb_parser *h = static_cast<dmlcb_parser *>(handle);
  return h->p64 != NULL ? h->p64->BytesRead() : h->p32->BytesRead();
}

void dmlcb_parser_free(void *handle) {
  dmlcb_parser *h = static_cast<dmlcb_parser *>(handle);
  if (h == NULL) return;
  delete h->p64;
  delete h->p32;
  delete h;
}

/*
 * Parser<I>::Create(uri, part, nparts, type) drained `passes` times with BeforeFirst between the
 * passes, touching only sizes -- the loop of the reference's throughput drivers
 * (test/libsvm_parser_test.cc:14-34, test/csv_parser_test.cc:22-57).  Returns the seconds the
 * passes took (Create not included), or -1 with a message.  hash7 (optional): FNV-1a of the
 * seven arrays over one extra, untimed pass, for comparing two implementations on one file.
 */
double dmlcb_parser_time(const char *uri, unsigned part, unsigned nparts, const char *type,
                         int index_bytes, int passes, uint64_t *rows, uint64_t *nnz,
                         uint64_t *bytes_read, uint64_t *hash7, double *create_seconds,
                         char *err, uint64_t errlen) {
  double seconds = -1.0;
  try {
    const double t0 = dmlc::GetTime();
    if (index_bytes == 8) {
      dmlc::Parser<uint64_t> *p = dmlc::Parser<uint64_t>::Create(uri, part, nparts, type);
      if (create_seconds != NULL) *create_seconds = dmlc::GetTime() - t0;
      try {
        DrainTimed(p, passes, hash7 != NULL, rows, nnz, hash7, &seconds);
        *bytes_read = p->BytesRead();
      } catch (...) {
        delete p;
        throw;
      }
      delete p;
    } else {
      dmlc::Parser<uint32_t> *p = dmlc::Parser<uint32_t>::Create(uri, part, nparts, type);
      if (create_seconds != NULL) *create_seconds = dmlc::GetTime() - t0;
      try {
        DrainTimed(p, passes, hash7 != NULL, rows, nnz, hash7, &seconds);
        *bytes_read = p->BytesRead();
      } catch (...) {
        delete p;
        throw;
      }
      delete p;
    }
  } catch (const std::exception &e) {
    Say(err, errlen, e.what());
    return -1.0;
  }
  return seconds;
}

/* RowBlockIter<I>::Create + NumCol (reference test/dataiter_test.cc); blocks come from
   dmlcb_rowiter_next and stay valid until the next call on the handle */
struct dmlcb_rowiter {
  dmlc::RowBlockIter<uint32_t> *i32;
  dmlc::RowBlockIter<uint64_t> *i64;
};

void *dmlcb_rowiter_create(const char *uri, unsigned part, unsigned nparts, const char *type,
                           int index_bytes, dmlcb_block *out, uint64_t *num_col, char *err,
                           uint64_t errlen) {
  dmlcb_rowiter *h = new dmlcb_rowiter();
  h->i32 = NULL;
  h->i64 = NULL;
  try {
    if (index_bytes == 8) {
      h->i64 = dmlc::RowBlockIter<uint64_t>::Create(uri, part, nparts, type);
      h->i64->BeforeFirst();
      *num_col = h->i64->NumCol();
    } else {
      h->i32 = dmlc::RowBlockIter<uint32_t>::Create(uri, part, nparts, type);
      h->i32->BeforeFirst();
      *num_col = h->i32->NumCol();
    }
  } catch (const std::exception &e) {
    Say(err, errlen, e.what());
    delete h->i32;
    delete h->i64;
    delete h;
    return NULL;
  }
  return h;
}

/* 1: another block was produced; 0: end */
int dmlcb_rowiter_next(void *handle, dmlcb_block *out) {
  dmlcb_rowiter *h = static_cast<dmlcb_rowiter *>(handle);
  if (h->i64 != NULL) {
    if (!h->i64->Next()) return 0;
    Export(h->i64->Value(), out);
  } else {
    if (!h->i32->Next()) return 0;
    Export(h->i32->Value(), out);
  }
  return 1;
}

void dmlcb_rowiter_before_first(void *handle) {
  dmlcb_rowiter *h = static_cast<dmlcb_rowiter *>(handle);
  if (h->i64 != NULL) h->i64->BeforeFirst(); else h->i32->BeforeFirst();
}

void dmlcb_rowiter_free(void *handle) {
  dmlcb_rowiter *h = static_cast<dmlcb_rowiter *>(handle);
  if (h == NULL) return;
  delete h->i32;
  delete h->i64;
  delete h;
}

/*
 * InputSplit::Create(uri, part, nparts, "text") drained with NextChunk (mode 0) or NextRecord
 * (mode 1, records re-joined with '\n'): the bytes of the partition, for checking the
 * partition arithmetic without a GPU. Returns the total length, or UINT64_MAX on error.
 */
uint64_t dmlcb_split_bytes(const char *uri, unsigned part, unsigned nparts, int mode,
                           uint64_t chunk_hint, char *buf, uint64_t cap, uint64_t *pieces,
                           char *err, uint64_t errlen) {
  uint64_t total = 0, n = 0;
  try {
    dmlc::InputSplit *split = dmlc::InputSplit::Create(uri, part, nparts, "text");
    if (chunk_hint != 0) split->HintChunkSize(chunk_hint);
    dmlc::InputSplit::Blob blob;
    while (mode == 0 ? split->NextChunk(&blob) : split->NextRecord(&blob)) {
      uint64_t len = blob.size;
      if (mode == 1) len = std::strlen(static_cast<const char *>(blob.dptr));
      if (total + len + 1 <= cap) {
        std::memcpy(buf + total, blob.dptr, len);
        if (mode == 1) buf[total + len] = '\n';
      }
      total += len + (mode == 1 ? 1 : 0);
      ++n;
    }
    delete split;
  } catch (const std::exception &e) {
    Say(err, errlen, e.what());
    return ~0ull;
  }
  if (pieces != NULL) *pieces = n;
  return total;
}

/*
 * InputSplitShuffle::Create(uri, part, nparts, "text", num_shuffle_parts, seed) drained with
 * NextChunk over `passes` passes (BeforeFirst between them): the bytes in the order they come
 * out, for comparing the sub-partition order with the reference's for a fixed seed.
 */
uint64_t dmlcb_shuffle_bytes(const char *uri, unsigned part, unsigned nparts, unsigned shuffle_parts,
                             int seed, int passes, char *buf, uint64_t cap, char *err, uint64_t errlen) {
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
    Say(err, errlen, e.what());
    return ~0ull;
  }
  return total;
}

}  // extern "C"
