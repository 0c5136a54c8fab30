"""ctypes binding of the C ABI in include/dmlc_b200_parser.h (libdmlc_b200.so).

This is the Python face of the product: the same entry points the C++ ``dmlc::Parser`` layer
calls.  It fails loudly when the CUDA library is missing or no GPU is usable -- there is no CPU
fallback anywhere in this package (the CPU oracle lives under oracle/ and is test-only).
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# DMLC_B200_LIB selects another build of the same library (the bounds-checked test build)
LIB_PATH = os.environ.get("DMLC_B200_LIB") or os.path.join(_HERE, "lib", "libdmlc_b200.so")

LIBSVM, LIBFM, CSV = 0, 1, 2
FORMATS = {"libsvm": LIBSVM, "libfm": LIBFM, "csv": CSV}
OK, E_INVALID, E_CUDA, E_PARSE, E_AGAIN, E_EMPTY, E_CAPACITY = 0, 1, 2, 3, 4, 5, 6
NOALLOC_MAX_BYTES = 256 << 20
MAX_CHUNK_BYTES = 1 << 30

# every symbol include/dmlc_b200_parser.h declares
EXPORTS = (
    "b200p_create", "b200p_destroy", "b200p_last_error", "b200p_parse_device",
    "b200p_parse_device_async", "b200p_parse_device_finish", "b200p_input_buffer",
    "b200p_submit", "b200p_next", "b200p_pending", "b200p_fetch", "b200p_launch_count",
    "b200p_set_host_threads", "b200p_copied_bytes", "b200p_next_to", "b200p_host_alloc",
    "b200p_host_free",
    "b200p_set_profiling", "b200p_kernel_time", "b200p_version", "b200p_parse_device_noalloc",
    "b200p_block_max", "b200p_current_device", "b200p_device_count", "b200p_set_delivery", "b200p_last_path",
)


class ParseError(RuntimeError):
    """The input makes the reference parser throw dmlc::Error (B200P_E_PARSE)."""


class Config(C.Structure):
    _fields_ = [("format", C.c_int32), ("index_bytes", C.c_int32), ("label_column", C.c_int32),
                ("device", C.c_int32), ("chunk_bytes", C.c_uint64), ("range_bytes", C.c_int32),
                ("depth", C.c_int32), ("stream", C.c_void_p), ("out_residency", C.c_int32),
                ("reserved", C.c_int32)]


class RowBlock(C.Structure):
    _fields_ = [("rows", C.c_uint64), ("nnz", C.c_uint64), ("n_weight", C.c_uint64),
                ("n_qid", C.c_uint64), ("n_field", C.c_uint64), ("n_value", C.c_uint64),
                ("offset", C.c_void_p), ("label", C.c_void_p), ("weight", C.c_void_p),
                ("qid", C.c_void_p), ("field", C.c_void_p), ("index", C.c_void_p),
                ("value", C.c_void_p), ("text_bytes", C.c_uint64), ("index_bytes", C.c_int32),
                ("on_device", C.c_int32)]


class Destination(C.Structure):
    """b200p_destination: caller-owned host arrays for b200p_next_to."""
    _fields_ = [("offset", C.c_void_p), ("label", C.c_void_p), ("weight", C.c_void_p),
                ("qid", C.c_void_p), ("field", C.c_void_p), ("index", C.c_void_p),
                ("value", C.c_void_p), ("cap_rows", C.c_uint64), ("cap_nnz", C.c_uint64),
                ("offset_base", C.c_uint64)]


class NoAllocRequest(C.Structure):
    _fields_ = [("format", C.c_int32), ("index_bytes", C.c_int32), ("label_column", C.c_int32),
                ("device", C.c_int32), ("stream", C.c_void_p), ("d_text", C.c_void_p),
                ("nbytes", C.c_uint64), ("d_offset", C.c_void_p), ("d_label", C.c_void_p),
                ("d_weight", C.c_void_p), ("d_qid", C.c_void_p), ("d_field", C.c_void_p),
                ("d_index", C.c_void_p), ("d_value", C.c_void_p), ("cap_rows", C.c_uint64),
                ("cap_nnz", C.c_uint64)]


_lib = None


def load():
    """Load libdmlc_b200.so (building nothing: see dmlc_core_b200.build)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("CUDA extension missing: %s (run `python -m dmlc_core_b200.build cuda`)"
                           % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    H = C.c_void_p
    lib.b200p_create.argtypes = [C.POINTER(Config), C.POINTER(H)]
    lib.b200p_create.restype = C.c_int
    lib.b200p_destroy.argtypes = [H]
    lib.b200p_destroy.restype = None
    lib.b200p_last_error.argtypes = [H]
    lib.b200p_last_error.restype = C.c_char_p
    lib.b200p_parse_device.argtypes = [H, C.c_void_p, C.c_uint64, C.POINTER(RowBlock)]
    lib.b200p_parse_device.restype = C.c_int
    lib.b200p_parse_device_async.argtypes = [H, C.c_void_p, C.c_uint64]
    lib.b200p_parse_device_async.restype = C.c_int
    lib.b200p_parse_device_finish.argtypes = [H, C.POINTER(RowBlock)]
    lib.b200p_parse_device_finish.restype = C.c_int
    lib.b200p_input_buffer.argtypes = [H, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
    lib.b200p_input_buffer.restype = C.c_int
    lib.b200p_submit.argtypes = [H, C.c_void_p, C.c_uint64]
    lib.b200p_submit.restype = C.c_int
    lib.b200p_next.argtypes = [H, C.POINTER(RowBlock)]
    lib.b200p_next.restype = C.c_int
    lib.b200p_pending.argtypes = [H]
    lib.b200p_pending.restype = C.c_int
    lib.b200p_fetch.argtypes = [H, C.c_void_p, C.c_void_p, C.c_uint64]
    lib.b200p_fetch.restype = C.c_int
    lib.b200p_next_to.argtypes = [H, C.POINTER(Destination), C.POINTER(RowBlock)]
    lib.b200p_next_to.restype = C.c_int
    lib.b200p_host_alloc.argtypes = [H, C.c_uint64, C.POINTER(C.c_void_p)]
    lib.b200p_host_alloc.restype = C.c_int
    lib.b200p_host_free.argtypes = [C.c_void_p]
    lib.b200p_host_free.restype = None
    lib.b200p_set_host_threads.argtypes = [H, C.c_int]
    lib.b200p_set_host_threads.restype = C.c_int
    lib.b200p_copied_bytes.argtypes = [H, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    lib.b200p_copied_bytes.restype = C.c_int
    lib.b200p_last_path.argtypes = [H]
    lib.b200p_last_path.restype = C.c_int
    lib.b200p_set_delivery.argtypes = [H, C.c_int]
    lib.b200p_set_delivery.restype = C.c_int
    lib.b200p_current_device.argtypes = []
    lib.b200p_current_device.restype = C.c_int
    lib.b200p_device_count.argtypes = []
    lib.b200p_device_count.restype = C.c_int
    lib.b200p_block_max.argtypes = [H, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    lib.b200p_block_max.restype = C.c_int
    lib.b200p_launch_count.argtypes = [H]
    lib.b200p_launch_count.restype = C.c_uint64
    lib.b200p_set_profiling.argtypes = [H, C.c_int]
    lib.b200p_set_profiling.restype = C.c_int
    lib.b200p_kernel_time.argtypes = [H, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_uint64)]
    lib.b200p_kernel_time.restype = C.c_int
    lib.b200p_parse_device_noalloc.argtypes = [C.POINTER(NoAllocRequest), C.POINTER(RowBlock),
                                               C.c_char_p, C.c_uint64]
    lib.b200p_parse_device_noalloc.restype = C.c_int
    lib.b200p_version.argtypes = []
    lib.b200p_version.restype = C.c_char_p
    _lib = lib
    return lib


def _host_array(ptr, n, dtype):
    if not ptr or n == 0:
        return np.zeros(0, dtype=dtype)
    buf = (C.c_char * (int(n) * np.dtype(dtype).itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype).copy()


class Parser(object):
    """One parser handle (one GPU, one format, one IndexType).

    Mirrors what a reference user holds after ``Parser<IndexType>::Create``: feed it chunks of
    text cut at line boundaries, get RowBlocks back.  ``parse_*`` return dicts with the seven
    RowBlock arrays as numpy arrays (offset, label, weight, qid, field, index, value).
    """

    def __init__(self, fmt, index_bytes=4, label_column=-1, device=0, chunk_bytes=0,
                 range_bytes=0, depth=0, stream=None, out_device=False):
        self.lib = load()
        cfg = Config()
        cfg.format = FORMATS.get(fmt, fmt)
        cfg.index_bytes = index_bytes
        cfg.label_column = label_column
        cfg.device = device
        cfg.chunk_bytes = chunk_bytes
        cfg.range_bytes = range_bytes
        cfg.depth = depth
        cfg.stream = stream
        cfg.out_residency = 1 if out_device else 0
        self.index_dtype = np.uint64 if index_bytes == 8 else np.uint32
        self.index_bytes = index_bytes
        self.handle = C.c_void_p()
        rc = self.lib.b200p_create(C.byref(cfg), C.byref(self.handle))
        if rc != OK:
            msg = self.lib.b200p_last_error(None).decode()
            self.handle = None
            raise RuntimeError("b200p_create failed (%d): %s" % (rc, msg))

    def close(self):
        if getattr(self, "handle", None):
            self.lib.b200p_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc == OK:
            return
        msg = self.lib.b200p_last_error(self.handle).decode()
        if rc == E_PARSE:
            raise ParseError(msg)
        raise RuntimeError("libdmlc_b200 error %d: %s" % (rc, msg))

    def set_delivery(self, on_demand):
        """b200p_set_delivery: False = arrays come home as soon as a chunk is parsed (default)."""
        self._check(self.lib.b200p_set_delivery(self.handle, 1 if on_demand else 0))

    # ---- device-resident path -------------------------------------------------------------
    def parse_device_async(self, d_ptr, nbytes):
        self._check(self.lib.b200p_parse_device_async(self.handle, d_ptr, nbytes))

    def parse_device_finish(self):
        blk = RowBlock()
        self._check(self.lib.b200p_parse_device_finish(self.handle, C.byref(blk)))
        return blk

    def parse_device(self, d_ptr, nbytes):
        """Parse a chunk resident in HBM; returns the raw RowBlock struct (device pointers)."""
        blk = RowBlock()
        self._check(self.lib.b200p_parse_device(self.handle, d_ptr, nbytes, C.byref(blk)))
        return blk

    def fetch(self, blk):
        """Copy a device-resident RowBlock to numpy arrays."""
        def take(ptr, n, dtype):
            out = np.zeros(int(n), dtype=dtype)
            if n and ptr:
                self._check(self.lib.b200p_fetch(self.handle, out.ctypes.data, ptr, out.nbytes))
            return out
        return {
            "offset": take(blk.offset, blk.rows + 1 if blk.rows else 0, np.uint64)
            if blk.rows else np.zeros(1, dtype=np.uint64),
            "label": take(blk.label, blk.rows, np.float32),
            "weight": take(blk.weight, blk.n_weight, np.float32),
            "qid": take(blk.qid, blk.n_qid, np.uint64),
            "field": take(blk.field, blk.n_field, self.index_dtype),
            "index": take(blk.index, blk.nnz, self.index_dtype),
            "value": take(blk.value, blk.n_value, np.float32),
            "error": 0,
        }

    # ---- streaming path -------------------------------------------------------------------
    def submit(self, data):
        """Enqueue a host chunk (bytes / bytearray / numpy uint8)."""
        arr = np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else data
        self._keep = arr
        self._check(self.lib.b200p_submit(self.handle, arr.ctypes.data, arr.nbytes))

    def input_buffer(self):
        ptr, cap = C.c_void_p(), C.c_uint64()
        self._check(self.lib.b200p_input_buffer(self.handle, C.byref(ptr), C.byref(cap)))
        return ptr.value, int(cap.value)

    def submit_in_place(self, ptr, nbytes):
        self._check(self.lib.b200p_submit(self.handle, ptr, nbytes))

    def next_raw(self):
        blk = RowBlock()
        self._check(self.lib.b200p_next(self.handle, C.byref(blk)))
        return blk

    def next(self):
        blk = self.next_raw()
        return self.to_numpy(blk)

    def to_numpy(self, blk):
        return {
            "offset": _host_array(blk.offset, blk.rows + 1, np.uint64) if blk.rows
            else np.zeros(1, dtype=np.uint64),
            "label": _host_array(blk.label, blk.rows, np.float32),
            "weight": _host_array(blk.weight, blk.n_weight, np.float32),
            "qid": _host_array(blk.qid, blk.n_qid, np.uint64),
            "field": _host_array(blk.field, blk.n_field, self.index_dtype),
            "index": _host_array(blk.index, blk.nnz, self.index_dtype),
            "value": _host_array(blk.value, blk.n_value, np.float32),
            "error": 0,
        }

    def block_max(self):
        """(max index, max field) of the block returned last, reduced on the GPU."""
        mi, mf = C.c_uint64(0), C.c_uint64(0)
        self._check(self.lib.b200p_block_max(self.handle, C.byref(mi), C.byref(mf)))
        return int(mi.value), int(mf.value)

    def pending(self):
        return int(self.lib.b200p_pending(self.handle))

    def set_profiling(self, on):
        self._check(self.lib.b200p_set_profiling(self.handle, 1 if on else 0))

    def kernel_time(self, kernel):
        """(total_ms, launches) of kernel 0=count, 1=scan, 2=emit since profiling was enabled."""
        ms, n = C.c_double(0), C.c_uint64(0)
        self._check(self.lib.b200p_kernel_time(self.handle, kernel, C.byref(ms), C.byref(n)))
        return float(ms.value), int(n.value)

    def launch_count(self):
        return int(self.lib.b200p_launch_count(self.handle))

    def last_path(self):
        """0 = two-pass walk, 1 = ParseTiles, 2 = CollectRanges / ConvertRanges (chunk finished last)."""
        return int(self.lib.b200p_last_path(self.handle))

    def host_alloc(self, nbytes):
        """Pinned, huge-page backed host memory (address); release with host_free()."""
        ptr = C.c_void_p()
        self._check(self.lib.b200p_host_alloc(self.handle, int(nbytes), C.byref(ptr)))
        return ptr.value

    def host_free(self, ptr):
        self.lib.b200p_host_free(ptr)

    def next_to(self, dest):
        """b200p_next_to: (rc, RowBlock); rc is E_CAPACITY (6) when the block stays queued."""
        blk = RowBlock()
        rc = self.lib.b200p_next_to(self.handle, C.byref(dest), C.byref(blk))
        if rc not in (0, 6):
            self._check(rc)
        return rc, blk

    def set_host_threads(self, threads):
        """Host threads that rebuild CSV column ids in next(); 0 = always copy them from the GPU."""
        self._check(self.lib.b200p_set_host_threads(self.handle, int(threads)))

    def copied_bytes(self):
        """(host-to-device, device-to-host) bytes this handle has moved over the link."""
        up, down = C.c_uint64(0), C.c_uint64(0)
        self._check(self.lib.b200p_copied_bytes(self.handle, C.byref(up), C.byref(down)))
        return int(up.value), int(down.value)

    def parse_host(self, data):
        """submit + next for one chunk; raises ParseError where the reference would throw."""
        self.submit(data)
        return self.next()


def parse_device_noalloc(request):
    """b200p_parse_device_noalloc: parse with caller-owned device arrays and the library's fixed
    scratch. Returns (rc, RowBlock, message); see include/dmlc_b200_parser.h for the limits."""
    lib = load()
    blk = RowBlock()
    err = C.create_string_buffer(512)
    rc = lib.b200p_parse_device_noalloc(C.byref(request), C.byref(blk), err, 512)
    return rc, blk, err.value.decode()
