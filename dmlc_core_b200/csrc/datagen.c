/*
 * Synthetic corpora for the five BASELINE.json configurations.
 * The reference ships no data files (its drivers take a user-supplied file,
 * reference test/libsvm_parser_test.cc:9-12), so tests and bench.py generate text here.
 *
 * Deterministic in (kind, seed, target_bytes) regardless of the thread count: rows are made in
 * blocks of kBlockRows, each block seeded by (seed, block id), and concatenated in block order.
 * Files always end with '\n', use single spaces and '\n' line ends, and contain no comments.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

enum {
  GEN_LIBSVM = 1,      /* cfg 1: label idx:val ..., nnz~U{20..40}, idx sorted distinct in [1,1e6) */
  GEN_CSV = 2,         /* cfg 2: label + 128 x %.6g U(-10,10), comma separated */
  GEN_LIBFM = 3,       /* cfg 3: label field:idx:val ..., field U{0..38}, idx U[0,2^40) */
  GEN_LIBSVM_QID = 4,  /* cfg 4: label:weight qid:Q idx idx ... (value-less features) */
  GEN_CRITEO = 5       /* cfg 5: label 13 dense k:int (30% missing) + 26 hash64:1 */
};

#define kBlockRows 512

typedef struct { uint64_t s; } rng_t;
static inline uint64_t rng_next(rng_t *r) { /* splitmix64 */
  uint64_t z = (r->s += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
static inline double rng_unit(rng_t *r) { return (double)(rng_next(r) >> 11) * (1.0 / 9007199254740992.0); }
static inline uint64_t rng_below(rng_t *r, uint64_t n) { return rng_next(r) % n; }

static int cmp_u64(const void *a, const void *b) {
  uint64_t x = *(const uint64_t *)a, y = *(const uint64_t *)b;
  return x < y ? -1 : x > y;
}

static char *put_u64(char *p, uint64_t v) {
  char tmp[24];
  int n = 0;
  do { tmp[n++] = (char)('0' + v % 10); v /= 10; } while (v);
  while (n) *p++ = tmp[--n];
  return p;
}

/* one block of rows into buf (capacity is generous: <= 2 KiB per row); returns bytes */
static size_t gen_block(int kind, uint64_t seed, uint64_t block, int ncols, char *buf) {
  rng_t r;
  r.s = seed * 0xD1342543DE82EF95ull + block * 0x2545F4914F6CDD1Dull + 12345;
  (void)rng_next(&r);
  char *p = buf;
  uint64_t qid = block * 1000u;
  int group_left = 0;
  for (int row = 0; row < kBlockRows; ++row) {
    if (kind == GEN_CSV) {
      *p++ = (char)('0' + (int)rng_below(&r, 2));
      for (int c = 0; c < ncols; ++c) {
        *p++ = ',';
        p += sprintf(p, "%.6g", rng_unit(&r) * 20.0 - 10.0);
      }
      *p++ = '\n';
      continue;
    }
    *p++ = (char)('0' + (int)rng_below(&r, 2));
    if (kind == GEN_LIBSVM_QID) {
      if (group_left == 0) { ++qid; group_left = 1 + (int)rng_below(&r, 50); }
      --group_left;
      *p++ = ':';
      p += sprintf(p, "%.6g", rng_unit(&r) * 2.0);
      memcpy(p, " qid:", 5); p += 5;
      p = put_u64(p, qid);
    }
    if (kind == GEN_CRITEO) {
      for (int k = 1; k <= 13; ++k) {
        if (rng_unit(&r) < 0.3) continue;
        *p++ = ' ';
        p = put_u64(p, (uint64_t)k);
        *p++ = ':';
        /* dense counts: heavy-tailed small integers */
        double u = rng_unit(&r);
        p = put_u64(p, (uint64_t)floor(exp(u * 9.0)) - 1u);
      }
      for (int k = 0; k < 26; ++k) {
        *p++ = ' ';
        p = put_u64(p, rng_next(&r));
        *p++ = ':'; *p++ = '1';
      }
      *p++ = '\n';
      continue;
    }
    int nnz = 20 + (int)rng_below(&r, 21);
    uint64_t idx[64];
    if (kind == GEN_LIBFM) {
      for (int k = 0; k < nnz; ++k) {
        *p++ = ' ';
        p = put_u64(p, rng_below(&r, 39));
        *p++ = ':';
        p = put_u64(p, rng_next(&r) & ((1ull << 40) - 1));
        *p++ = ':';
        p += sprintf(p, "%.6g", rng_unit(&r));
      }
    } else {
      for (int k = 0; k < nnz; ++k) idx[k] = 1 + rng_below(&r, 999999);
      qsort(idx, (size_t)nnz, sizeof(uint64_t), cmp_u64);
      for (int k = 1; k < nnz; ++k) if (idx[k] <= idx[k - 1]) idx[k] = idx[k - 1] + 1;
      for (int k = 0; k < nnz; ++k) {
        *p++ = ' ';
        p = put_u64(p, idx[k]);
        if (kind == GEN_LIBSVM) {
          *p++ = ':';
          p += sprintf(p, "%.6g", rng_unit(&r));
        }
      }
    }
    *p++ = '\n';
  }
  return (size_t)(p - buf);
}

/*
 * Fill `out` with whole blocks until at least target_bytes are produced (or cap would be
 * exceeded).  ncols: CSV feature columns (0 -> 128).  Returns bytes written; *rows_out rows.
 */
uint64_t b200gen_generate(int kind, uint64_t seed, uint64_t target_bytes, int ncols, char *out,
                          uint64_t cap, uint64_t *rows_out) {
  if (ncols <= 0) ncols = 128;
  const size_t slot = (size_t)kBlockRows * (kind == GEN_CSV ? (size_t)ncols * 16 + 64 : 2048);
  int nthreads = 1;
#ifdef _OPENMP
  nthreads = omp_get_max_threads();
#endif
  const int wave = nthreads * 4;
  char *scratch = (char *)malloc(slot * (size_t)wave);
  size_t *sizes = (size_t *)malloc(sizeof(size_t) * (size_t)wave);
  uint64_t written = 0, rows = 0, block = 0;
  int full = 0;
  while (!full && written < target_bytes) {
#pragma omp parallel for schedule(dynamic, 1)
    for (int i = 0; i < wave; ++i) {
      sizes[i] = gen_block(kind, seed, block + (uint64_t)i, ncols, scratch + slot * (size_t)i);
    }
    for (int i = 0; i < wave; ++i) {
      if (written >= target_bytes) break;
      if (written + sizes[i] > cap) { full = 1; break; }
      memcpy(out + written, scratch + slot * (size_t)i, sizes[i]);
      written += sizes[i];
      rows += kBlockRows;
    }
    block += (uint64_t)wave;
  }
  free(scratch);
  free(sizes);
  if (rows_out) *rows_out = rows;
  return written;
}
