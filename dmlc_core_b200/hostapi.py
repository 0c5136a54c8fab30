"""ctypes binding of the C++ host layer (libdmlc_b200_host.so): the dmlc::Parser<IndexType> /
RowBlockIter / InputSplit API of this package, driven the way the reference's drivers drive
theirs (Create -> Next()/Value() loop -> BeforeFirst -> delete)."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libdmlc_b200_host.so")
_lib = None


class DmlcError(RuntimeError):
    """dmlc::Error (or dmlc::ParamError) raised by the C++ layer."""


class Block(C.Structure):
    _fields_ = [("size", C.c_uint64), ("offset", C.c_void_p), ("label", C.c_void_p),
                ("weight", C.c_void_p), ("qid", C.c_void_p), ("field", C.c_void_p),
                ("index", C.c_void_p), ("value", C.c_void_p)]


def load(path=None):
    """Load the host library. `path` selects another build of the same C wrapper (the tests use
    it for the variant compiled against the reference's own headers)."""
    global _lib
    if path is None and _lib is not None:
        return _lib
    lib_path = path or LIB_PATH
    if not os.path.exists(lib_path):
        raise RuntimeError("host library missing: %s (python -m dmlc_core_b200.build host)" % lib_path)
    from . import capi
    capi.load()  # libdmlc_b200.so first (rpath $ORIGIN also finds it)
    lib = C.CDLL(lib_path)
    lib.dmlcb_parser_create.argtypes = [C.c_char_p, C.c_uint, C.c_uint, C.c_char_p, C.c_int,
                                        C.c_char_p, C.c_uint64]
    lib.dmlcb_parser_create.restype = C.c_void_p
    lib.dmlcb_parser_next.argtypes = [C.c_void_p, C.POINTER(Block), C.c_char_p, C.c_uint64]
    lib.dmlcb_parser_next.restype = C.c_int
    lib.dmlcb_parser_before_first.argtypes = [C.c_void_p]
    lib.dmlcb_parser_before_first.restype = None
    lib.dmlcb_parser_bytes_read.argtypes = [C.c_void_p]
    lib.dmlcb_parser_bytes_read.restype = C.c_uint64
    lib.dmlcb_parser_free.argtypes = [C.c_void_p]
    lib.dmlcb_parser_free.restype = None
    lib.dmlcb_parser_time.argtypes = [C.c_char_p, C.c_uint, C.c_uint, C.c_char_p, C.c_int, C.c_int,
                                      C.POINTER(C.c_uint64), C.POINTER(C.c_uint64),
                                      C.POINTER(C.c_uint64), C.POINTER(C.c_uint64),
                                      C.POINTER(C.c_double), C.c_char_p, C.c_uint64]
    lib.dmlcb_parser_time.restype = C.c_double
    lib.dmlcb_rowiter_create.argtypes = [C.c_char_p, C.c_uint, C.c_uint, C.c_char_p, C.c_int,
                                         C.POINTER(Block), C.POINTER(C.c_uint64), C.c_char_p,
                                         C.c_uint64]
    lib.dmlcb_rowiter_create.restype = C.c_void_p
    lib.dmlcb_rowiter_next.argtypes = [C.c_void_p, C.POINTER(Block)]
    lib.dmlcb_rowiter_next.restype = C.c_int
    lib.dmlcb_rowiter_before_first.argtypes = [C.c_void_p]
    lib.dmlcb_rowiter_before_first.restype = None
    lib.dmlcb_rowiter_free.argtypes = [C.c_void_p]
    lib.dmlcb_rowiter_free.restype = None
    lib.dmlcb_split_bytes.argtypes = [C.c_char_p, C.c_uint, C.c_uint, C.c_int, C.c_uint64,
                                      C.c_char_p, C.c_uint64, C.POINTER(C.c_uint64), C.c_char_p,
                                      C.c_uint64]
    lib.dmlcb_split_bytes.restype = C.c_uint64
    if path is None:
        _lib = lib
    return lib


def _arr(ptr, n, dtype):
    if not ptr or n == 0:
        return np.zeros(0, dtype=dtype)
    buf = (C.c_char * (int(n) * np.dtype(dtype).itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype).copy()


def _block_to_numpy(b, index_bytes):
    idt = np.uint64 if index_bytes == 8 else np.uint32
    n = int(b.size)
    offset = _arr(b.offset, n + 1, np.uint64)
    nnz = int(offset[-1] - offset[0]) if n else 0
    base = int(offset[0]) if n else 0
    isz = np.dtype(idt).itemsize
    return {
        "offset": offset - np.uint64(base),
        "label": _arr(b.label, n, np.float32),
        "weight": _arr(b.weight, n, np.float32),
        "qid": _arr(b.qid, n, np.uint64),
        "field": _arr(b.field + base * isz if b.field else None, nnz if b.field else 0, idt),
        "index": _arr(b.index + base * isz if b.index else None, nnz if b.index else 0, idt),
        "value": _arr(b.value + base * 4 if b.value else None, nnz if b.value else 0, np.float32),
    }


def time_parser(uri, part=0, nparts=1, type_="libsvm", index_bytes=4, passes=1, want_hash=False, lib=None):
    """Parser<I>::Create(uri, part, nparts, type) drained `passes` times (BeforeFirst between
    them), sizes only: the reference's throughput driver loop. Returns a dict with seconds
    (Create excluded), create_seconds, rows / nnz (summed over the passes), bytes_read and, if
    asked, the FNV-1a hashes of the seven arrays from one extra untimed pass."""
    lib = lib if lib is not None else load()
    rows, nnz, nbytes = C.c_uint64(0), C.c_uint64(0), C.c_uint64(0)
    h = (C.c_uint64 * 7)()
    created = C.c_double(0)
    err = C.create_string_buffer(2048)
    sec = lib.dmlcb_parser_time(uri.encode(), part, nparts, type_.encode(), index_bytes, passes,
                                C.byref(rows), C.byref(nnz), C.byref(nbytes),
                                h if want_hash else None, C.byref(created), err, 2048)
    if sec < 0:
        raise DmlcError(err.value.decode())
    return {"seconds": float(sec), "create_seconds": float(created.value), "rows": int(rows.value),
            "nnz": int(nnz.value), "bytes_read": int(nbytes.value),
            "hash": [int(x) for x in h] if want_hash else None}


class CppParser(object):
    """dmlc::Parser<IndexType>::Create(uri, part_index, num_parts, type)."""

    def __init__(self, uri, part_index=0, num_parts=1, type_="libsvm", index_bytes=4, lib=None):
        self.lib = lib if lib is not None else load()
        self.index_bytes = index_bytes
        err = C.create_string_buffer(2048)
        self.h = self.lib.dmlcb_parser_create(uri.encode(), part_index, num_parts, type_.encode(),
                                              index_bytes, err, 2048)
        if not self.h:
            raise DmlcError(err.value.decode())

    def next(self):
        """One RowBlock as numpy arrays, or None at the end."""
        b = Block()
        err = C.create_string_buffer(2048)
        rc = self.lib.dmlcb_parser_next(self.h, C.byref(b), err, 2048)
        if rc < 0:
            raise DmlcError(err.value.decode())
        if rc == 0:
            return None
        return _block_to_numpy(b, self.index_bytes)

    def before_first(self):
        self.lib.dmlcb_parser_before_first(self.h)

    def bytes_read(self):
        return int(self.lib.dmlcb_parser_bytes_read(self.h))

    def close(self):
        if self.h:
            self.lib.dmlcb_parser_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def drain(self):
        """Concatenation over all blocks, offset rebased to a running nnz total (the parity
        contract of SURVEY.md 8a). Returns (dict, number of blocks)."""
        parts, nblocks = [], 0
        while True:
            b = self.next()
            if b is None:
                break
            parts.append(b)
            nblocks += 1
        return concat_blocks(parts, self.index_bytes), nblocks


def concat_blocks(parts, index_bytes):
    idt = np.uint64 if index_bytes == 8 else np.uint32
    out = {}
    for k, dt in (("label", np.float32), ("weight", np.float32), ("qid", np.uint64), ("field", idt),
                  ("index", idt), ("value", np.float32)):
        out[k] = np.concatenate([p[k] for p in parts]) if parts else np.zeros(0, dtype=dt)
    offs, run = [np.zeros(1, dtype=np.uint64)], 0
    for p in parts:
        offs.append(p["offset"][1:] + np.uint64(run))
        run += len(p["index"])
    out["offset"] = np.concatenate(offs)
    out["error"] = 0
    return out


def row_iter(uri, part_index=0, num_parts=1, type_="libsvm", index_bytes=4, lib=None, passes=1):
    """RowBlockIter<IndexType>::Create(...) drained `passes` times (BeforeFirst in between):
    returns (concatenated arrays of the last pass, NumCol(), number of blocks per pass)."""
    lib = lib if lib is not None else load()
    b, ncol = Block(), C.c_uint64(0)
    err = C.create_string_buffer(2048)
    h = lib.dmlcb_rowiter_create(uri.encode(), part_index, num_parts, type_.encode(), index_bytes,
                                 C.byref(b), C.byref(ncol), err, 2048)
    if not h:
        raise DmlcError(err.value.decode())
    parts = []
    for k in range(passes):
        if k:
            lib.dmlcb_rowiter_before_first(h)
        parts = []
        while lib.dmlcb_rowiter_next(h, C.byref(b)) == 1:
            parts.append(_block_to_numpy(b, index_bytes))
    out = concat_blocks(parts, index_bytes)
    lib.dmlcb_rowiter_free(h)
    return out, int(ncol.value), len(parts)


def split_bytes(uri, part_index, num_parts, mode=0, chunk_hint=0, cap=1 << 26):
    """Bytes InputSplit::Create(uri, part, nparts, "text") yields via NextChunk (mode 0) or
    NextRecord (mode 1). Returns (bytes, pieces)."""
    lib = load()
    buf = C.create_string_buffer(cap + 1)
    pieces = C.c_uint64(0)
    err = C.create_string_buffer(2048)
    n = lib.dmlcb_split_bytes(uri.encode(), part_index, num_parts, mode, chunk_hint, buf, cap,
                              C.byref(pieces), err, 2048)
    if n == 2 ** 64 - 1:
        raise DmlcError(err.value.decode())
    return buf.raw[:n], int(pieces.value)
