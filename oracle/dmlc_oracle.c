/*
 * TEST INFRASTRUCTURE ONLY -- the CPU checker for the CUDA parsers.
 *
 * A plain-C restatement of the text -> CSR hot path of rahul003/dmlc-core, written from the
 * behaviour of the cited reference lines, not copied from them.  Only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
 * this library; the product (dmlc_core_b200/) never does and has no CPU fallback.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks every function here against
 *   (1) the 14 literals of the reference's own test/strtonum_test.cc:8-10,18-20,28-30,
 *   (2) the bit patterns of SURVEY.md Appendix A,
 *   (3) tests/golden/ -- inputs plus RowBlocks produced by the compiled, unmodified reference
 *       (oracle/_ref, generator script oracle/make_golden.py), and
 *   (4) oracle/_ref itself on randomized corpora whenever that library is present.
 *
 * Conventions: a "source" is a byte buffer with a hard limit; reads at or past the limit give
 * NUL, which is what the reference sees after a chunk (input_split_base.cc:247 zero-pads).
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "parse_result.h"

typedef struct {
  const unsigned char *b;
  size_t lim;
} src_t;

static inline unsigned at(const src_t *s, size_t i) { return i < s->lim ? s->b[i] : 0u; }

/* character classes: reference src/data/strtonum.h:14-31 */
static inline int is_space(unsigned c) { return c == ' ' || c == '\t' || c == '\r' || c == '\n' || c == '\f'; }
static inline int is_blank(unsigned c) { return c == ' ' || c == '\t'; }
static inline int is_digit(unsigned c) { return c >= '0' && c <= '9'; }
static inline int is_digitchar(unsigned c) {
  return is_digit(c) || c == '+' || c == '-' || c == '.' || c == 'e' || c == 'E';
}
static inline int is_eol(unsigned c) { return c == '\n' || c == '\r'; }

/*
 * Float conversion: reference src/data/strtonum.h:37-97.
 *   - integer digits accumulate in fp32, one rounding per multiply and per add (:52-54);
 *   - fraction digits accumulate in two wrapping uint64 (digits, power of ten), divided in
 *     fp64, rounded to fp32, added in fp32 (:57-68);
 *   - exponent digits accumulate in a wrapping `unsigned`, clamp to 38, scale built in an fp32
 *     variable by fp64 multiplies of 1e8 then 10 (:71-93).
 * Compiled with -ffp-contract=off so no step is fused.
 */
static float conv_float(const src_t *s, size_t p, size_t *endp) {
  while (is_space(at(s, p))) ++p;
  int negative = 0;
  if (at(s, p) == '-') {
    negative = 1;
    ++p;
  } else if (at(s, p) == '+') {
    ++p;
  }
  volatile float acc = 0.0f; /* volatile: forbid excess precision / reassociation */
  while (is_digit(at(s, p))) {
    float t = acc * 10.0f;
    acc = t + (float)(int)(at(s, p) - '0');
    ++p;
  }
  if (at(s, p) == '.') {
    uint64_t digits = 0, scale10 = 1;
    ++p;
    while (is_digit(at(s, p))) {
      digits = digits * 10u + (uint64_t)(at(s, p) - '0');
      scale10 *= 10u;
      ++p;
    }
    double q = (double)digits / (double)scale10;
    acc = acc + (float)q;
  }
  if (at(s, p) == 'e' || at(s, p) == 'E') {
    ++p;
    int divide = 0;
    if (at(s, p) == '-') {
      divide = 1;
      ++p;
    } else if (at(s, p) == '+') {
      ++p;
    }
    uint32_t ex = 0;
    while (is_digit(at(s, p))) {
      ex = ex * 10u + (uint32_t)(at(s, p) - '0');
      ++p;
    }
    if (ex > 38u) ex = 38u;
    volatile float sc = 1.0f;
    while (ex >= 8u) {
      sc = (float)((double)sc * 1E8);
      ex -= 8u;
    }
    while (ex > 0u) {
      sc = (float)((double)sc * 10.0);
      ex -= 1u;
    }
    acc = divide ? (acc / sc) : (acc * sc);
  }
  if (endp != NULL) *endp = p;
  return negative ? -acc : acc;
}

/*
 * Unsigned index conversion: reference src/data/strtonum.h:126-150 reached through
 * Str2T<uint32_t> (:186-190, accumulates in a 32-bit int => value mod 2^32) or
 * Str2T<uint64_t> (:202-207, mod 2^64).  A '-' sign makes the reference CHECK fail (:141).
 */
static uint64_t conv_uint(const src_t *s, size_t p, int width, int *neg_error) {
  while (is_space(at(s, p))) ++p;
  if (at(s, p) == '-') {
    *neg_error = 1;
    return 0;
  }
  if (at(s, p) == '+') ++p;
  uint64_t v = 0;
  while (is_digit(at(s, p))) {
    v = v * 10u + (uint64_t)(at(s, p) - '0');
    ++p;
  }
  return width == 4 ? (uint64_t)(uint32_t)v : v;
}

/*
 * glibc atoll as used for `qid:` (reference src/data/libsvm_parser.h:69-75): skip C-locale
 * whitespace (space \t \n \v \f \r), optional sign, decimal digits, saturate at
 * LLONG_MAX / LLONG_MIN; the caller casts to uint64_t.
 */
static int64_t conv_atoll(const src_t *s, size_t p) {
  for (;;) {
    unsigned c = at(s, p);
    if (c == ' ' || (c >= '\t' && c <= '\r')) {
      ++p;
    } else {
      break;
    }
  }
  int negative = 0;
  if (at(s, p) == '-') {
    negative = 1;
    ++p;
  } else if (at(s, p) == '+') {
    ++p;
  }
  const uint64_t cap = negative ? (uint64_t)1 << 63 : ((uint64_t)1 << 63) - 1;
  uint64_t v = 0;
  int saturated = 0;
  while (is_digit(at(s, p))) {
    unsigned d = at(s, p) - '0';
    if (!saturated) {
      if (v > (cap - d) / 10u) {
        saturated = 1;
        v = cap;
      } else {
        v = v * 10u + d;
      }
    }
    ++p;
  }
  if (negative) return (int64_t)(0u - v);
  return (int64_t)v;
}

/* growable output --------------------------------------------------------------------- */
typedef struct {
  uint64_t *offset; size_t n_offset, c_offset;
  float *label; size_t n_label, c_label;
  float *weight; size_t n_weight, c_weight;
  uint64_t *qid; size_t n_qid, c_qid;
  uint64_t *field; size_t n_field, c_field; /* kept 64-bit, narrowed on export */
  uint64_t *index; size_t n_index, c_index;
  float *value; size_t n_value, c_value;
  int neg_error;
} acc_t;

#define PUSH(a, name, type, v)                                                        \
  do {                                                                                \
    if ((a)->n_##name == (a)->c_##name) {                                             \
      (a)->c_##name = (a)->c_##name ? (a)->c_##name * 2 : 1024;                       \
      (a)->name = (type *)realloc((a)->name, (a)->c_##name * sizeof(type));           \
    }                                                                                 \
    (a)->name[(a)->n_##name++] = (v);                                                 \
  } while (0)

/*
 * One "v1[:v2[:v3]]" group: reference ParsePair / ParseTriple, src/data/strtonum.h:227-303.
 * Starting at p inside [p, lend): skip non-digitchars; take the digitchar run as v1; then up to
 * (arity-1) times: skip blanks, and if a ':' follows, skip every non-digitchar and take the next
 * run.  starts[k] receives the start of run k (conversion starts there and ignores the run end,
 * strtonum.h:211-216).  Returns how many runs were bound (0..arity); *endp is where the
 * reference leaves its cursor.
 */
static int scan_group(const src_t *s, size_t p, size_t lend, int arity, size_t starts[3],
                      size_t *endp) {
  while (p != lend && !is_digitchar(at(s, p))) ++p;
  if (p == lend) {
    *endp = lend;
    return 0;
  }
  int got = 0;
  for (;;) {
    starts[got++] = p;
    size_t q = p;
    while (q != lend && is_digitchar(at(s, q))) ++q;
    if (got == arity) {
      *endp = q;
      return got;
    }
    p = q;
    while (p != lend && is_blank(at(s, p))) ++p;
    if (p == lend || at(s, p) != ':') {
      *endp = p;
      return got;
    }
    ++p;
    while (p != lend && !is_digitchar(at(s, p))) ++p;
    /* p may equal lend here: the reference then converts "at lend" (an empty run) */
  }
}

/* first EOL strictly after lbegin, or end: reference libsvm_parser.h:44-46 (the first byte of
   a line is never tested) */
static size_t line_end(const src_t *s, size_t lbegin, size_t end) {
  size_t e = lbegin + 1;
  while (e != end && !is_eol(at(s, e))) ++e;
  return e;
}

/* LibSVM block: reference src/data/libsvm_parser.h:35-99 */
static void block_libsvm(const src_t *s, size_t begin, size_t end, int width, acc_t *a) {
  size_t lbegin = begin;
  while (lbegin != end) {
    size_t lend = line_end(s, lbegin, end);
    size_t st[3], q;
    int r = scan_group(s, lbegin, lend, 2, st, &q);
    if (r < 1) {
      lbegin = lend;
      continue;
    }
    float label = conv_float(s, st[0], NULL);
    if (r == 2) PUSH(a, weight, float, conv_float(s, st[1], NULL));
    if (a->n_label != 0) PUSH(a, offset, uint64_t, (uint64_t)a->n_index);
    PUSH(a, label, float, label);
    size_t p = q;
    while (p != end && at(s, p) == ' ') ++p;
    if (p != lend && at(s, p) == 'q' && at(s, p + 1) == 'i' && at(s, p + 2) == 'd' &&
        at(s, p + 3) == ':') {
      PUSH(a, qid, uint64_t, (uint64_t)conv_atoll(s, p + 4));
    }
    p = q; /* the reference rewinds, so "qid:N" is parsed again as a feature (:74) */
    while (p != lend) {
      r = scan_group(s, p, lend, 2, st, &q);
      if (r >= 1) {
        PUSH(a, index, uint64_t, conv_uint(s, st[0], width, &a->neg_error));
        if (r == 2) PUSH(a, value, float, conv_float(s, st[1], NULL));
      }
      p = q;
    }
    lbegin = lend;
  }
  if (a->n_label != 0) PUSH(a, offset, uint64_t, (uint64_t)a->n_index);
}

/* LibFM block: reference src/data/libfm_parser.h:35-93 */
static void block_libfm(const src_t *s, size_t begin, size_t end, int width, acc_t *a) {
  size_t lbegin = begin;
  while (lbegin != end) {
    size_t lend = line_end(s, lbegin, end);
    size_t st[3], q;
    int r = scan_group(s, lbegin, lend, 2, st, &q);
    if (r < 1) {
      lbegin = lend;
      continue;
    }
    float label = conv_float(s, st[0], NULL);
    if (r == 2) PUSH(a, weight, float, conv_float(s, st[1], NULL));
    if (a->n_label != 0) PUSH(a, offset, uint64_t, (uint64_t)a->n_index);
    PUSH(a, label, float, label);
    size_t p = q;
    while (p != lend) {
      r = scan_group(s, p, lend, 3, st, &q);
      if (r >= 2) {
        PUSH(a, field, uint64_t, conv_uint(s, st[0], width, &a->neg_error));
        PUSH(a, index, uint64_t, conv_uint(s, st[1], width, &a->neg_error));
        if (r == 3) PUSH(a, value, float, conv_float(s, st[2], NULL));
      } else if (r == 1) {
        /* a lone token is dropped, but the reference has already converted it as a field id,
           so a '-' sign still raises (libfm_parser.h:73 -> strtonum.h:141) */
        (void)conv_uint(s, st[0], width, &a->neg_error);
      }
      p = q;
    }
    lbegin = lend;
  }
  if (a->n_label != 0) PUSH(a, offset, uint64_t, (uint64_t)a->n_index);
}

/* CSV block: reference src/data/csv_parser.h:63-104 */
static void block_csv(const src_t *s, size_t begin, size_t end, int label_column, acc_t *a) {
  size_t lbegin = begin;
  while (lbegin != end && is_eol(at(s, lbegin))) ++lbegin;
  while (lbegin != end) {
    size_t lend = line_end(s, lbegin, end);
    size_t p = lbegin;
    int column = 0;
    uint64_t idx = 0;
    float label = 0.0f;
    while (p != lend) {
      size_t e;
      float v = conv_float(s, p, &e);
      p = e;
      if (column == label_column) {
        label = v;
      } else {
        PUSH(a, value, float, v);
        PUSH(a, index, uint64_t, idx);
        ++idx;
      }
      ++column;
      while (at(s, p) != ',' && p != lend) ++p;
      if (p != lend) ++p;
    }
    while (lend != end && is_eol(at(s, lend))) ++lend;
    lbegin = lend;
    PUSH(a, label, float, label);
    PUSH(a, offset, uint64_t, (uint64_t)a->n_index);
  }
}

static void *narrow(const uint64_t *v, size_t n, int width) {
  if (n == 0) return NULL;
  if (width == 8) {
    uint64_t *o = (uint64_t *)malloc(n * 8);
    memcpy(o, v, n * 8);
    return o;
  }
  uint32_t *o = (uint32_t *)malloc(n * 4);
  for (size_t i = 0; i < n; ++i) o[i] = (uint32_t)v[i];
  return o;
}

static void *shrink(void *p, size_t n, size_t elt) {
  if (n == 0) {
    free(p);
    return NULL;
  }
  return realloc(p, n * elt);
}

/*
 * Parse one block held in memory.  Consistency checks are the reference's:
 * src/data/row_block.h:173-177 (GetBlock) and src/data/libfm_parser.h:91.
 */
void oracle_parse_buffer(int format, int index_bytes, int label_column, const char *buf,
                         size_t len, oracle_result *out) {
  memset(out, 0, sizeof(*out));
  acc_t a;
  memset(&a, 0, sizeof(a));
  PUSH(&a, offset, uint64_t, 0);
  src_t s = {(const unsigned char *)buf, len};
  const int width = index_bytes == 8 ? 8 : 4;
  if (format == ORACLE_LIBSVM) {
    block_libsvm(&s, 0, len, width, &a);
  } else if (format == ORACLE_LIBFM) {
    block_libfm(&s, 0, len, width, &a);
  } else {
    block_csv(&s, 0, len, label_column, &a);
  }
  if (a.neg_error) {
    out->error = 1;
    snprintf(out->errmsg, sizeof(out->errmsg), "Check failed: sign == true (negative unsigned field)");
  } else if (a.n_value != 0 && a.n_value != a.n_index) {
    out->error = 1;
    snprintf(out->errmsg, sizeof(out->errmsg),
             "Check failed: offset.back() == value.size() || value.size() == 0");
  } else if (format == ORACLE_LIBFM && a.n_field != a.n_index) {
    out->error = 1;
    snprintf(out->errmsg, sizeof(out->errmsg), "Check failed: field.size() == index.size()");
  }
  out->index_bytes = width;
  out->bytes_read = len;
  out->n_blocks = a.n_label != 0 ? 1 : 0;
  out->n_offset = a.n_offset;
  out->n_label = a.n_label;
  out->n_weight = a.n_weight;
  out->n_qid = a.n_qid;
  out->n_field = a.n_field;
  out->n_index = a.n_index;
  out->n_value = a.n_value;
  out->offset = (uint64_t *)shrink(a.offset, a.n_offset, 8);
  out->label = (float *)shrink(a.label, a.n_label, 4);
  out->weight = (float *)shrink(a.weight, a.n_weight, 4);
  out->qid = (uint64_t *)shrink(a.qid, a.n_qid, 8);
  out->field = narrow(a.field, a.n_field, width);
  out->index = narrow(a.index, a.n_index, width);
  out->value = (float *)shrink(a.value, a.n_value, 4);
  free(a.field);
  free(a.index);
}

void oracle_free(oracle_result *r) {
  free(r->offset);
  free(r->label);
  free(r->weight);
  free(r->qid);
  free(r->field);
  free(r->index);
  free(r->value);
  memset(r, 0, sizeof(*r));
}

/* s is NUL terminated (the limit is its length). */
uint32_t oracle_strtof_bits(const char *str, int64_t *consumed) {
  src_t s = {(const unsigned char *)str, strlen(str)};
  size_t e = 0;
  float v = conv_float(&s, 0, &e);
  if (consumed != NULL) *consumed = (int64_t)e;
  uint32_t bits;
  memcpy(&bits, &v, 4);
  return bits;
}

uint64_t oracle_strtouint(const char *str, int width, int *threw) {
  src_t s = {(const unsigned char *)str, strlen(str)};
  *threw = 0;
  return conv_uint(&s, 0, width, threw);
}

int64_t oracle_atoll(const char *str) {
  src_t s = {(const unsigned char *)str, strlen(str)};
  return conv_atoll(&s, 0);
}

/*
 * Partition arithmetic: reference src/io/input_split_base.cc:30-64 with align_bytes == 1
 * (text) and src/io/line_split.cc:9-26 (SeekRecordBegin).  `sizes` are the sizes of the files
 * that make up the logical byte stream, in listing order; read_at(ctx, file, offset) returns
 * the byte there or -1 at that file's end.  Outputs the adjusted [begin, end) in stream offsets.
 */
typedef int (*oracle_read_fn)(void *ctx, size_t file, uint64_t off);

static uint64_t seek_record_begin(oracle_read_fn rd, void *ctx, size_t file, uint64_t off) {
  uint64_t n = 0;
  for (;;) { /* up to and including the first EOL */
    int c = rd(ctx, file, off + n);
    if (c < 0) return n;
    ++n;
    if (c == '\n' || c == '\r') break;
  }
  for (;;) { /* plus every EOL byte that follows immediately */
    int c = rd(ctx, file, off + n);
    if (c < 0) return n;
    if (c != '\n' && c != '\r') break;
    ++n;
  }
  return n;
}

void oracle_partition(const uint64_t *sizes, size_t nfiles, unsigned rank, unsigned nsplit,
                      oracle_read_fn rd, void *ctx, uint64_t *begin, uint64_t *end) {
  uint64_t total = 0;
  for (size_t i = 0; i < nfiles; ++i) total += sizes[i];
  uint64_t step = (total + nsplit - 1) / nsplit;
  uint64_t b = step * rank < total ? step * rank : total;
  uint64_t e = step * (rank + 1) < total ? step * (rank + 1) : total;
  *begin = b;
  *end = e;
  if (b == e) return;
  /* file holding an offset = last file whose start is <= offset (upper_bound - 1, :38-43) */
  uint64_t start = 0;
  size_t fb = 0, fe = 0;
  uint64_t fb_start = 0, fe_start = 0;
  for (size_t i = 0; i < nfiles; ++i) {
    if (start <= b) { fb = i; fb_start = start; }
    if (start <= e) { fe = i; fe_start = start; }
    start += sizes[i];
  }
  if (start <= e) { fe = nfiles; fe_start = start; } /* e == total: one past the last file */
  if (e != fe_start) *end = e + seek_record_begin(rd, ctx, fe, e - fe_start);
  if (b != fb_start) *begin = b + seek_record_begin(rd, ctx, fb, b - fb_start);
}

/* Convenience over files on disk: concatenated bytes of partition (rank, nsplit). */
typedef struct {
  FILE **fp;
} files_ctx;

static int file_read_at(void *ctx, size_t file, uint64_t off) {
  files_ctx *f = (files_ctx *)ctx;
  if (fseeko(f->fp[file], (off_t)off, SEEK_SET) != 0) return -1;
  int c = fgetc(f->fp[file]);
  return c == EOF ? -1 : c;
}

uint64_t oracle_split_bytes(const char **paths, size_t nfiles, unsigned rank, unsigned nsplit,
                            char *buf, uint64_t cap, uint64_t *begin_out, uint64_t *end_out) {
  uint64_t *sizes = (uint64_t *)calloc(nfiles, 8);
  files_ctx ctx;
  ctx.fp = (FILE **)calloc(nfiles, sizeof(FILE *));
  for (size_t i = 0; i < nfiles; ++i) {
    ctx.fp[i] = fopen(paths[i], "rb");
    if (ctx.fp[i] == NULL) { sizes[i] = 0; continue; }
    fseeko(ctx.fp[i], 0, SEEK_END);
    sizes[i] = (uint64_t)ftello(ctx.fp[i]);
  }
  uint64_t b, e;
  oracle_partition(sizes, nfiles, rank, nsplit, file_read_at, &ctx, &b, &e);
  if (begin_out) *begin_out = b;
  if (end_out) *end_out = e;
  uint64_t written = 0, start = 0;
  for (size_t i = 0; i < nfiles; ++i) {
    uint64_t lo = b > start ? b : start;
    uint64_t hi = e < start + sizes[i] ? e : start + sizes[i];
    if (lo < hi && ctx.fp[i] != NULL) {
      uint64_t want = hi - lo;
      if (written + want > cap) want = cap > written ? cap - written : 0;
      fseeko(ctx.fp[i], (off_t)(lo - start), SEEK_SET);
      if (want != 0 && fread(buf + written, 1, want, ctx.fp[i]) != want) { /* short read */ }
      written += hi - lo;
    }
    start += sizes[i];
  }
  for (size_t i = 0; i < nfiles; ++i) if (ctx.fp[i]) fclose(ctx.fp[i]);
  free(ctx.fp);
  free(sizes);
  return e > b ? e - b : 0;
}
