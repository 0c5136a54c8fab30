/*
 * dmlc_b200_parser.h -- C ABI of the B200 text -> CSR parser (libdmlc_b200.so).
 *
 * This is the whole boundary between a host program (the dmlc::Parser<IndexType> shim in
 * dmlc_core_b200/src, a ctypes binding, cgo, JNI ...) and the sm_100a kernels: plain pointers
 * and sizes, no C++ or torch types.  Each entry point names the reference interface it stands
 * in for (paths relative to rahul003/dmlc-core).
 *
 * Threading: a handle is single-consumer, like a dmlc::Parser instance
 * (reference include/dmlc/data.h:283-311). Distinct handles (e.g. one per GPU / part_index) are
 * independent.
 *
 * Return values: 0 = success; B200P_E_* otherwise. After a failure b200p_last_error(handle)
 * describes it. B200P_E_PARSE is the data error the reference raises as dmlc::Error
 * (negative unsigned field: src/data/strtonum.h:141; value/index count mismatch:
 * src/data/row_block.h:177) -- the C++ layer turns it back into `throw dmlc::Error`.
 */
#ifndef DMLC_B200_PARSER_H_
#define DMLC_B200_PARSER_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200P_OK 0
#define B200P_E_INVALID 1   /* bad argument / call sequence */
#define B200P_E_CUDA 2      /* CUDA runtime failure (no usable GPU, out of memory ...) */
#define B200P_E_PARSE 3     /* input makes the reference parser throw dmlc::Error */
#define B200P_E_AGAIN 4     /* no free pipeline slot: call b200p_next first */
#define B200P_E_EMPTY 5     /* nothing submitted */
#define B200P_E_CAPACITY 6  /* caller-provided output arrays are too small (no-alloc entry only) */

#define B200P_FORMAT_LIBSVM 0 /* src/data/libsvm_parser.h */
#define B200P_FORMAT_LIBFM 1  /* src/data/libfm_parser.h */
#define B200P_FORMAT_CSV 2    /* src/data/csv_parser.h */

/* largest chunk one call may carry (token positions are 31-bit on the device) */
#define B200P_MAX_CHUNK_BYTES ((uint64_t)1 << 30)

typedef struct b200p_parser_s *b200p_handle;

/*
 * Parser configuration.  Stands in for the constructor arguments of
 * LibSVMParser / LibFMParser / CSVParser (src/data/libsvm_parser.h:24-27,
 * src/data/csv_parser.h:43-52) plus the template IndexType of dmlc::Parser.
 */
typedef struct b200p_config {
  int32_t format;        /* B200P_FORMAT_* */
  int32_t index_bytes;   /* sizeof(IndexType): 4 or 8 */
  int32_t label_column;  /* CSVParserParam::label_column (csv_parser.h:24,31), -1 = none */
  int32_t device;        /* CUDA device ordinal */
  uint64_t chunk_bytes;  /* expected chunk size, sizes the staging buffers (0 = 64 MiB) */
  int32_t range_bytes;   /* bytes of text one warp owns (0 = default); tuning only */
  int32_t depth;         /* chunks in flight for b200p_submit/b200p_next (0 = 3) */
  void *stream;          /* cudaStream_t for the device-resident calls; NULL = own stream */
  int32_t out_residency; /* b200p_submit / b200p_next: B200P_OUT_HOST (0, default) returns the arrays
                            in pinned host memory; B200P_OUT_DEVICE leaves them in HBM for a consumer
                            that lives on the GPU (no device-to-host copy at all) */
  int32_t reserved;      /* 0 */
} b200p_config;

#define B200P_OUT_HOST 0
#define B200P_OUT_DEVICE 1

/*
 * One parsed chunk.  Field for field the RowBlock<IndexType> of include/dmlc/data.h:169-231
 * (size, offset, label, weight, qid, field, index, value), plus the true lengths of the
 * optional arrays, which the reference hides behind raw pointers.  Arrays that stayed empty
 * are NULL, as RowBlockContainer::GetBlock does with BeginPtr (src/data/row_block.h:180-186).
 * offset[0] == 0; offset has rows + 1 entries.
 */
typedef struct b200p_rowblock {
  uint64_t rows;       /* RowBlock::size */
  uint64_t nnz;        /* offset[rows] */
  uint64_t n_weight;   /* entries in weight (0 or, for well-formed input, rows) */
  uint64_t n_qid;
  uint64_t n_field;    /* 0 or nnz */
  uint64_t n_value;    /* 0 or nnz */
  const uint64_t *offset;
  const float *label;
  const float *weight;
  const uint64_t *qid;
  const void *field;   /* index_bytes wide */
  const void *index;   /* index_bytes wide */
  const float *value;
  uint64_t text_bytes; /* size of the chunk this block came from (-> Parser::BytesRead) */
  int32_t index_bytes;
  int32_t on_device;   /* 1: pointers are device memory, 0: pinned host memory */
} b200p_rowblock;

/* Replaces: new LibSVMParser<I>(source, nthread) etc. (src/data.cc:21-60). Allocates the
   device-side staging; fails with B200P_E_CUDA when there is no GPU -- there is no CPU path. */
int b200p_create(const b200p_config *config, b200p_handle *out);

/* Replaces: delete parser (src/data/text_parser.h:37-39). */
void b200p_destroy(b200p_handle h);

/* The CUDA device current on the calling thread (-1 without a usable GPU) and the number of
   devices: lets a host layer that does not link the CUDA runtime default to "the GPU my caller
   selected". No reference counterpart. */
int b200p_current_device(void);
int b200p_device_count(void);

/* Text of the last failure on this handle ("" if none). h may be NULL for create failures. */
const char *b200p_last_error(b200p_handle h);

/*
 * Device-resident parse: the chunk already sits in HBM (16-byte aligned, nbytes <=
 * B200P_MAX_CHUNK_BYTES, cut after an end-of-line byte unless it is the last of its stream) and
 * the CSR arrays stay in HBM (out->on_device == 1, valid until the next call on this handle).
 * Replaces: TextParserBase::FillData + ParseBlock on one chunk (src/data/text_parser.h:93-134).
 * b200p_parse_device = b200p_parse_device_async + b200p_parse_device_finish.
 * The async half only enqueues the kernels on the handle's stream and never synchronises.
 * One chunk per finish: a second b200p_parse_device_async before the finish returns
 * B200P_E_AGAIN (use several handles, one stream each, to keep several chunks in flight).
 */
int b200p_parse_device(b200p_handle h, const void *d_text, uint64_t nbytes, b200p_rowblock *out);
int b200p_parse_device_async(b200p_handle h, const void *d_text, uint64_t nbytes);
int b200p_parse_device_finish(b200p_handle h, b200p_rowblock *out);

/*
 * Streaming parse from host memory: H2D copy, kernels and D2H copy of chunk k overlap with
 * those of chunks k+1.. on separate streams (the GPU-side counterpart of the reference's two
 * prefetch stages, src/io/threaded_input_split.h:29-42 and src/data/parser.h:73-82).
 *
 *   b200p_input_buffer  pinned staging buffer of the next free slot, to be filled in place
 *                       (e.g. by pread); capacity >= config.chunk_bytes.
 *   b200p_submit        enqueue a chunk. `text` may be the pointer b200p_input_buffer returned
 *                       or memory the caller pinned itself (both copied to the device from
 *                       where they lie; caller-pinned memory must stay untouched until the
 *                       matching b200p_next), or pageable memory (staged through the pinned
 *                       buffer first).
 *   b200p_next          wait for the oldest submitted chunk; pointers are pinned host memory
 *                       (device memory for B200P_OUT_DEVICE handles, out->on_device == 1),
 *                       valid until the next b200p_next / b200p_destroy on this handle.
 * Replaces: ParserImpl::ParseNext -> FillData (src/data/parser.h:30-44, text_parser.h:93-134).
 *
 * Threads: a handle may be driven by two threads at once -- one feeding (b200p_input_buffer,
 * b200p_submit) and one draining (b200p_next, b200p_pending, b200p_block_max, b200p_last_error)
 * -- which is how the C++ Parser keeps reading files while the consumer holds a block (the
 * reference's ThreadedInputSplit / ThreadedParser pairing). Anything beyond that pairing, and
 * every other entry point, needs the caller's own serialisation.
 */
int b200p_input_buffer(b200p_handle h, void **host_ptr, uint64_t *capacity);
int b200p_submit(b200p_handle h, const void *text, uint64_t nbytes);
int b200p_next(b200p_handle h, b200p_rowblock *out);
/* chunks submitted and not yet returned by b200p_next */
int b200p_pending(b200p_handle h);

/*
 * When the arrays of a chunk travel to the host (B200P_OUT_HOST handles).
 *   B200P_DELIVER_EAGER      (default) as soon as the chunk's kernels have finished, into the
 *                            slot's own pinned arrays, by a thread inside the library: the copies
 *                            run while the caller is still working on the previous block, and
 *                            b200p_next only waits for them.  This is the reference's ThreadedParser
 *                            running ahead of its consumer (src/data/parser.h:73-82).
 *   B200P_DELIVER_ON_DEMAND  when b200p_next / b200p_next_to is called: for b200p_next_to users,
 *                            whose destination is not known before the call (a block that was
 *                            already brought home eagerly is copied on from the slot's arrays).
 */
#define B200P_DELIVER_EAGER 0
#define B200P_DELIVER_ON_DEMAND 1
int b200p_set_delivery(b200p_handle h, int mode);

/*
 * b200p_next with the arrays written straight into memory of the caller's choosing, at the
 * caller's running position: what RowBlockContainer::Push(RowBlock) does for BasicRowIter
 * (src/data/row_block.h:122-166, src/data/basic_row_iter.h:62-82) without the intermediate block
 * and its memcpy. offset receives rows + 1 entries, each increased by offset_base (so that the
 * first one overwrites the previous block's last entry with the same value); the other arrays
 * receive exactly the counts reported in *out. Absent arrays may be NULL. Memory from
 * b200p_host_alloc is pinned and takes the copies at link speed; any other host memory works,
 * slower. If the block does not fit (cap_rows / cap_nnz), B200P_E_CAPACITY is returned, out->rows
 * and out->nnz say what it needs, and the block stays queued for the next call.
 */
typedef struct b200p_destination {
  uint64_t *offset;      /* cap_rows + 1 entries */
  float *label;          /* cap_rows */
  float *weight;         /* cap_rows (LibSVM, LibFM) */
  uint64_t *qid;         /* cap_rows (LibSVM) */
  void *field;           /* cap_nnz entries of index_bytes (LibFM) */
  void *index;           /* cap_nnz entries of index_bytes */
  float *value;          /* cap_nnz */
  uint64_t cap_rows;
  uint64_t cap_nnz;
  uint64_t offset_base;
} b200p_destination;
int b200p_next_to(b200p_handle h, const b200p_destination *dst, b200p_rowblock *out);

/* Pinned host memory on the handle's device, huge-page backed where the host allows it (what the
   library uses for its own staging and result buffers). b200p_host_free works after the handle
   is gone. No reference counterpart. */
int b200p_host_alloc(b200p_handle h, uint64_t bytes, void **ptr);
void b200p_host_free(void *ptr);

/*
 * Host threads b200p_next may use (0..64). Their one job: for CSV chunks whose rows average 16
 * or more entries, index[] is not copied back from the GPU -- it is implied by offset[] (the j-th
 * non-label field of a row has column id j, src/data/csv_parser.h:84-88) -- but written into the
 * pinned result by these threads while value[] is still on the link; that halves the
 * device-to-host bytes of the path. The same threads widen LibFM field ids, which cross the
 * link as one byte each when every id of the chunk is below 256. 0 = always copy the full arrays
 * from the GPU. Default: a quarter of
 * the host's cores (divided by LOCAL_WORLD_SIZE when a launcher exports it), at most 4; the
 * environment variable B200P_HOST_THREADS overrides the default.
 * The arrays returned are identical either way. No reference counterpart.
 */
int b200p_set_host_threads(b200p_handle h, int threads);

/* bytes this handle has moved over the host link so far: text in (b200p_submit) and arrays out
   (b200p_next, b200p_fetch). bench.py's e2e byte counts. No reference counterpart. */
int b200p_copied_bytes(b200p_handle h, uint64_t *host_to_device, uint64_t *device_to_host);

/*
 * Largest value in index[] and field[] (0 when empty) of the block most recently returned by
 * b200p_next or b200p_parse_device[_finish] on this handle, reduced on the GPU from the arrays
 * still resident there. Replaces the per-element max_index / max_field bookkeeping of
 * RowBlockContainer::Push (src/data/row_block.h:143,153) behind RowBlockIter::NumCol()
 * (src/data/basic_row_iter.h:44). On demand, because a streaming Parser user never needs it.
 */
int b200p_block_max(b200p_handle h, uint64_t *max_index, uint64_t *max_field);

/* Copy nbytes from device memory returned by the device-resident calls into host memory,
   ordered after the handle's stream (for callers that do not link the CUDA runtime).
   No reference counterpart: the reference never leaves host memory. */
int b200p_fetch(b200p_handle h, void *host_dst, const void *device_src, uint64_t nbytes);

/* number of kernels this handle has launched so far (bench.py's gpu_launches claim) */
uint64_t b200p_launch_count(b200p_handle h);

/*
 * Which kernels parsed the chunk finished last on this handle (-1 before the first):
 *   B200P_PATH_COLLECT_CONVERT  CollectRanges, ScanTiles, ScanFinish, ConvertRanges: LibSVM and LibFM
 *   B200P_PATH_ONE_LAUNCH       ParseTiles: CSV with the label in column 0 or no label column
 *   B200P_PATH_TWO_PASS         WalkText (count), the scan, WalkText (emit): CSV with a later label
 *                               column, and every chunk the other two hand over (a line longer than
 *                               their windows, text denser than their queues)
 * All three produce the same arrays. No reference counterpart.
 */
#define B200P_PATH_TWO_PASS 0
#define B200P_PATH_ONE_LAUNCH 1
#define B200P_PATH_COLLECT_CONVERT 2
int b200p_last_path(b200p_handle h);

/*
 * Allocation-free variant of b200p_parse_device. It has no counterpart in the reference.
 *
 * For callers that cannot allocate GPU memory while they run: the card was handed over in full
 * to a memory pool at start-up (RMM under XGBoost, a framework's caching allocator), or the
 * process is in the middle of NCCL traffic where cudaMalloc/cudaFree stalls the device. Such a
 * caller passes the text AND every CSR array (device pointers from its own pool, capacities in
 * elements); the kernels' scratch sits in a fixed __device__ array inside the library, about
 * 640 KiB that each GPU loading the library keeps until exit. No handle is needed and the call
 * performs no cudaMalloc / cudaFree / pinned allocation.
 *
 * Limit: nbytes <= B200P_NOALLOC_MAX_BYTES (256 MiB); larger chunks return B200P_E_INVALID with
 * a message in errbuf. Arrays too small: B200P_E_CAPACITY, out->rows / out->nnz = what the chunk
 * needs, nothing written. Required arrays: offset (cap_rows + 1 entries), label, index, value;
 * weight unless CSV; qid for LibSVM; field for LibFM.
 *
 * The scratch is shared by all calls on a GPU and keeps the previous call's contents; the call
 * clears what it needs cleared, so back-to-back calls are fine, but calls that overlap on the
 * same GPU (two streams, two threads) race -- serialise them. The call synchronises `stream`
 * before returning. The first call on each device looks up the scratch's device address
 * (cudaGetSymbolAddress); make that first call before starting any stream capture.
 */
#define B200P_NOALLOC_MAX_BYTES ((uint64_t)256 << 20)
typedef struct b200p_noalloc_request {
  int32_t format;        /* B200P_FORMAT_* */
  int32_t index_bytes;   /* 4 or 8 */
  int32_t label_column;  /* CSV */
  int32_t device;        /* CUDA device ordinal the pointers below belong to */
  void *stream;          /* cudaStream_t to run on (NULL = default stream) */
  const void *d_text;    /* chunk in device memory, 16-byte aligned, cut after a line end */
  uint64_t nbytes;
  uint64_t *d_offset;    /* outputs, all device memory owned by the caller */
  float *d_label;
  float *d_weight;
  uint64_t *d_qid;
  void *d_field;
  void *d_index;
  float *d_value;
  uint64_t cap_rows;     /* rows the row-sized arrays hold (d_offset holds cap_rows + 1) */
  uint64_t cap_nnz;      /* entries d_field / d_index / d_value hold */
} b200p_noalloc_request;
int b200p_parse_device_noalloc(const b200p_noalloc_request *request, b200p_rowblock *out,
                               char *errbuf, uint64_t errlen);

/*
 * Per-kernel device time, measured with CUDA events recorded on the stream each kernel is
 * launched on. Off by default (two extra event records per launch). b200p_kernel_time waits for
 * the recorded launches and returns their summed duration and count since profiling was
 * switched on. No reference counterpart (the reference only prints wall-clock MB/s,
 * test/libsvm_parser_test.cc:27-31).
 */
#define B200P_KERNEL_COUNT 0 /* CollectRanges, or pass 1 of the walk (WalkText<EMIT=false>) */
#define B200P_KERNEL_SCAN 1  /* ScanTiles + ScanFinish */
#define B200P_KERNEL_EMIT 2  /* ConvertRanges, ParseTiles, or pass 2 of the walk (WalkText<EMIT=true>) */
int b200p_set_profiling(b200p_handle h, int on);
int b200p_kernel_time(b200p_handle h, int kernel, double *total_ms, uint64_t *launches);

/* library / build identification: "dmlc_b200 <version> sm_100a" */
const char *b200p_version(void);

#ifdef __cplusplus
}
#endif
#endif /* DMLC_B200_PARSER_H_ */
