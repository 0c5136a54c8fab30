/*
 * TEST INFRASTRUCTURE ONLY.  Shared result layout used by the two CPU checkers
 * under oracle/ (the compiled reference shim in oracle/_ref and the C restatement
 * oracle/dmlc_oracle.c).  Nothing under dmlc_core_b200/ includes this file.
 *
 * The layout is the parity contract of SURVEY.md section 8(a): the concatenation,
 * over every RowBlock a parser returns, of label / weight / qid / field / index /
 * value, with `offset` rebased to a running nnz total (offset[0] == 0).
 */
#ifndef ORACLE_PARSE_RESULT_H_
#define ORACLE_PARSE_RESULT_H_
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct oracle_result {
  uint64_t n_offset;  /* rows + 1 (1 when there are no rows) */
  uint64_t n_label;   /* rows */
  uint64_t n_weight;
  uint64_t n_qid;
  uint64_t n_field;
  uint64_t n_index;   /* nnz */
  uint64_t n_value;
  uint64_t *offset;
  float *label;
  float *weight;
  uint64_t *qid;
  void *field;        /* index_bytes wide */
  void *index;        /* index_bytes wide */
  float *value;
  int32_t index_bytes;
  int32_t error;      /* non-zero: the parser raised dmlc::Error (message below) */
  uint64_t bytes_read;
  uint64_t n_blocks;
  char errmsg[512];
} oracle_result;

enum { ORACLE_LIBSVM = 0, ORACLE_LIBFM = 1, ORACLE_CSV = 2 };

#ifdef __cplusplus
}
#endif
#endif  /* ORACLE_PARSE_RESULT_H_ */
