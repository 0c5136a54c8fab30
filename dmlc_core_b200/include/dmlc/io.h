/*
 * The slice of reference include/dmlc/io.h the parser path touches: the InputSplit interface
 * (io.h:135-244).  InputSplit::Create is implemented in dmlc_core_b200/src/text_split.cc for
 * type "text" on local files with the reference's partition semantics
 * (src/io/input_split_base.cc:30-64, src/io/line_split.cc:9-34).
 */
#ifndef DMLC_IO_H_
#define DMLC_IO_H_

#include <cstddef>

#include "./base.h"
#include "./logging.h"

namespace dmlc {

/* a record-aligned view of one partition of a set of input files */
class InputSplit {
 public:
  /* a piece of memory owned by the split, valid until the next call on it */
  struct Blob {
    void *dptr;
    size_t size;
  };
  virtual ~InputSplit(void) {}
  /* advisory size for NextChunk (io.h:152) */
  virtual void HintChunkSize(size_t chunk_size) {}
  /* total bytes of all files behind the URI */
  virtual size_t GetTotalSize(void) = 0;
  /* rewind to the first byte of the partition */
  virtual void BeforeFirst(void) = 0;
  /* next line (without its end-of-line bytes, NUL terminated); false at the end */
  virtual bool NextRecord(Blob *out_rec) = 0;
  /* next run of whole lines; false at the end */
  virtual bool NextChunk(Blob *out_chunk) = 0;
  virtual bool NextBatch(Blob *out_chunk, size_t n_records) { return NextChunk(out_chunk); }
  /* re-aim this split at another partition of the same files */
  virtual void ResetPartition(unsigned part_index, unsigned num_parts) = 0;
  /*
   * uri: ';'-separated files and/or directories, or "stdin"; "uri#cachefile" records the chunks of
   * the first pass in cachefile and replays them afterwards; type: "text".
   * Part part_index of num_parts = bytes [ceil(T/n)*k, ceil(T/n)*(k+1)) of the concatenated
   * files, each end moved forward past the next end-of-line run unless it is a file start.
   */
  static InputSplit *Create(const char *uri, unsigned part_index, unsigned num_parts,
                            const char *type);
};

}  // namespace dmlc
#endif  // DMLC_IO_H_
