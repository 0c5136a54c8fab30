"""TEST INFRASTRUCTURE ONLY: ctypes access to the two CPU checkers.

* ``Oracle()``   -> oracle/_build/liboracle.so, the C restatement (oracle/dmlc_oracle.c)
* ``Reference()``-> oracle/_ref/libdmlc_ref.so, the unmodified reference behind oracle/ref_shim.cc
                    (``None`` when that library was never built / did not travel)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import
this module.  The product package (dmlc_core_b200) must never import it.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIBSVM, LIBFM, CSV = 0, 1, 2
FORMATS = {"libsvm": LIBSVM, "libfm": LIBFM, "csv": CSV}


class _Result(C.Structure):
    _fields_ = [
        ("n_offset", C.c_uint64), ("n_label", C.c_uint64), ("n_weight", C.c_uint64),
        ("n_qid", C.c_uint64), ("n_field", C.c_uint64), ("n_index", C.c_uint64),
        ("n_value", C.c_uint64),
        ("offset", C.c_void_p), ("label", C.c_void_p), ("weight", C.c_void_p),
        ("qid", C.c_void_p), ("field", C.c_void_p), ("index", C.c_void_p),
        ("value", C.c_void_p),
        ("index_bytes", C.c_int32), ("error", C.c_int32),
        ("bytes_read", C.c_uint64), ("n_blocks", C.c_uint64),
        ("errmsg", C.c_char * 512),
    ]


def _take(ptr, n, dtype):
    if not ptr or n == 0:
        return np.zeros(0, dtype=dtype)
    buf = (C.c_char * (int(n) * np.dtype(dtype).itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype).copy()


def _to_dict(r):
    idt = np.uint64 if r.index_bytes == 8 else np.uint32
    return {
        "offset": _take(r.offset, r.n_offset, np.uint64),
        "label": _take(r.label, r.n_label, np.float32),
        "weight": _take(r.weight, r.n_weight, np.float32),
        "qid": _take(r.qid, r.n_qid, np.uint64),
        "field": _take(r.field, r.n_field, idt),
        "index": _take(r.index, r.n_index, idt),
        "value": _take(r.value, r.n_value, np.float32),
        "error": int(r.error),
        "errmsg": r.errmsg.decode("utf-8", "replace"),
        "bytes_read": int(r.bytes_read),
        "n_blocks": int(r.n_blocks),
    }


ARRAYS = ("offset", "label", "weight", "qid", "field", "index", "value")


def same_rowblocks(a, b):
    """Bit-exact comparison of two result dicts (floats compared as their IEEE bits).
    Returns the name of the first array that differs, or None."""
    if bool(a["error"]) != bool(b["error"]):
        return "error"
    if a["error"]:
        return None
    for k in ARRAYS:
        x, y = a[k], b[k]
        if x.dtype != y.dtype or x.shape != y.shape:
            return k
        if x.dtype == np.float32:
            x, y = x.view(np.uint32), y.view(np.uint32)
        if not np.array_equal(x, y):
            return k
    return None


class _Checker(object):
    prefix = None

    def __init__(self, lib):
        self.lib = lib
        p = self.prefix
        self._parse_buffer = getattr(lib, p + "_parse_buffer")
        self._parse_buffer.argtypes = [C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_size_t,
                                       C.POINTER(_Result)]
        self._parse_buffer.restype = None
        self._free = getattr(lib, p + "_free")
        self._free.argtypes = [C.POINTER(_Result)]
        self._free.restype = None
        self._strtof = getattr(lib, p + "_strtof_bits")
        self._strtof.argtypes = [C.c_char_p, C.POINTER(C.c_int64)]
        self._strtof.restype = C.c_uint32
        self._strtouint = getattr(lib, p + "_strtouint")
        self._strtouint.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_int)]
        self._strtouint.restype = C.c_uint64

    def parse_buffer(self, fmt, data, index_bytes=4, label_column=-1):
        fmt = FORMATS.get(fmt, fmt)
        r = _Result()
        data = bytes(data)
        self._parse_buffer(fmt, index_bytes, label_column, data, len(data), C.byref(r))
        out = _to_dict(r)
        self._free(C.byref(r))
        return out

    def strtof_bits(self, token):
        n = C.c_int64(0)
        if isinstance(token, str):
            token = token.encode()
        bits = self._strtof(token, C.byref(n))
        return int(bits), int(n.value)

    def strtouint(self, token, width):
        t = C.c_int(0)
        if isinstance(token, str):
            token = token.encode()
        v = self._strtouint(token, width, C.byref(t))
        return int(v), bool(t.value)


class Oracle(_Checker):
    """The C restatement. Built on demand with oracle/Makefile (gcc only)."""
    prefix = "oracle"

    def __init__(self):
        path = os.path.join(HERE, "_build", "liboracle.so")
        src = os.path.join(HERE, "dmlc_oracle.c")
        if (not os.path.exists(path)) or os.path.getmtime(path) < os.path.getmtime(src):
            subprocess.check_call(["make", "-s", "-C", HERE, "oracle"])
        lib = C.CDLL(path)
        super(Oracle, self).__init__(lib)
        lib.oracle_atoll.argtypes = [C.c_char_p]
        lib.oracle_atoll.restype = C.c_int64
        lib.oracle_split_bytes.argtypes = [C.POINTER(C.c_char_p), C.c_size_t, C.c_uint, C.c_uint,
                                           C.c_char_p, C.c_uint64, C.POINTER(C.c_uint64),
                                           C.POINTER(C.c_uint64)]
        lib.oracle_split_bytes.restype = C.c_uint64

    def atoll(self, token):
        if isinstance(token, str):
            token = token.encode()
        return int(self.lib.oracle_atoll(token))

    def split_bytes(self, paths, part, nparts):
        """(bytes of the partition, begin, end) for a list of files, per the reference's
        ResetPartition arithmetic."""
        arr = (C.c_char_p * len(paths))(*[p.encode() for p in paths])
        total = sum(os.path.getsize(p) for p in paths)
        buf = C.create_string_buffer(total + 1)
        b, e = C.c_uint64(0), C.c_uint64(0)
        n = self.lib.oracle_split_bytes(arr, len(paths), part, nparts, buf, total, C.byref(b),
                                        C.byref(e))
        return buf.raw[:n], int(b.value), int(e.value)

    def parse_files(self, fmt, paths, part=0, nparts=1, index_bytes=4, label_column=-1):
        data, _, _ = self.split_bytes(paths, part, nparts)
        out = self.parse_buffer(fmt, data, index_bytes, label_column)
        return out


class Reference(_Checker):
    """The compiled, unmodified reference (oracle/_ref). Use Reference.load() -> None if absent."""
    prefix = "refshim"

    @staticmethod
    def path():
        return os.path.join(HERE, "_ref", "libdmlc_ref.so")

    @classmethod
    def load(cls):
        if not os.path.exists(cls.path()):
            return None
        try:
            return cls()
        except OSError:
            return None

    def __init__(self):
        lib = C.CDLL(self.path())
        super(Reference, self).__init__(lib)
        lib.refshim_parse_uri.argtypes = [C.c_char_p, C.c_uint, C.c_uint, C.c_char_p, C.c_int,
                                          C.POINTER(_Result)]
        lib.refshim_parse_uri.restype = None
        lib.refshim_split_bytes.argtypes = [C.c_char_p, C.c_uint, C.c_uint, C.c_char_p,
                                            C.c_uint64, C.POINTER(C.c_uint64)]
        lib.refshim_split_bytes.restype = C.c_uint64
        lib.refshim_time_uri.argtypes = [C.c_char_p, C.c_uint, C.c_uint, C.c_char_p, C.c_int,
                                         C.c_int, C.c_int, C.POINTER(C.c_uint64),
                                         C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        lib.refshim_time_uri.restype = C.c_double
        lib.refshim_max_threads.restype = C.c_int
        lib.refshim_iter_uri.argtypes = [C.c_char_p, C.c_uint, C.c_uint, C.c_char_p, C.c_int,
                                         C.POINTER(_Result), C.POINTER(C.c_uint64)]
        lib.refshim_iter_uri.restype = None

    def parse_uri(self, uri, part=0, nparts=1, type_="libsvm", index_bytes=4):
        r = _Result()
        self.lib.refshim_parse_uri(uri.encode(), part, nparts, type_.encode(), index_bytes,
                                   C.byref(r))
        out = _to_dict(r)
        self._free(C.byref(r))
        return out

    def iter_uri(self, uri, part=0, nparts=1, type_="libsvm", index_bytes=4):
        """RowBlockIter<I>::Create drained once -> (arrays, NumCol)."""
        r = _Result()
        ncol = C.c_uint64(0)
        self.lib.refshim_iter_uri(uri.encode(), part, nparts, type_.encode(), index_bytes,
                                  C.byref(r), C.byref(ncol))
        out = _to_dict(r)
        self._free(C.byref(r))
        return out, int(ncol.value)

    def split_bytes(self, uri, part, nparts, cap):
        buf = C.create_string_buffer(cap + 1)
        nch = C.c_uint64(0)
        n = self.lib.refshim_split_bytes(uri.encode(), part, nparts, buf, cap, C.byref(nch))
        return buf.raw[:min(n, cap)], int(n), int(nch.value)

    def time_uri(self, uri, part=0, nparts=1, type_="libsvm", index_bytes=4, nthread=0,
                 passes=1):
        rows, nnz, nbytes = C.c_uint64(0), C.c_uint64(0), C.c_uint64(0)
        sec = self.lib.refshim_time_uri(uri.encode(), part, nparts, type_.encode(), index_bytes,
                                        nthread, passes, C.byref(rows), C.byref(nnz),
                                        C.byref(nbytes))
        return {"seconds": float(sec), "rows": int(rows.value), "nnz": int(nnz.value),
                "bytes": int(nbytes.value)}

    def hash_uri(self, uri, part=0, nparts=1, type_="libsvm", index_bytes=4, nthread=0):
        """FNV-1a of the seven arrays (row lengths for offset) of one drain -> (list of 7, rows, nnz)."""
        h = (C.c_uint64 * 7)()
        rows, nnz = C.c_uint64(0), C.c_uint64(0)
        self.lib.refshim_hash_uri.argtypes = [C.c_char_p, C.c_uint, C.c_uint, C.c_char_p, C.c_int, C.c_int,
                                              C.POINTER(C.c_uint64), C.POINTER(C.c_uint64),
                                              C.POINTER(C.c_uint64)]
        self.lib.refshim_hash_uri.restype = C.c_int
        rc = self.lib.refshim_hash_uri(uri.encode(), part, nparts, type_.encode(), index_bytes, nthread,
                                       h, C.byref(rows), C.byref(nnz))
        if rc != 0:
            raise RuntimeError("refshim_hash_uri failed")
        return [int(x) for x in h], int(rows.value), int(nnz.value)

    def max_threads(self):
        return int(self.lib.refshim_max_threads())
