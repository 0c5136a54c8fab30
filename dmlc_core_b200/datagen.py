"""Synthetic corpora for the BASELINE.json configurations (csrc/datagen.c via ctypes)."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libb200gen.so")

LIBSVM, CSVK, LIBFM, LIBSVM_QID, CRITEO = 1, 2, 3, 4, 5
KINDS = {"libsvm": LIBSVM, "csv": CSVK, "libfm": LIBFM, "libsvm_qid": LIBSVM_QID,
         "criteo": CRITEO}
_lib = None


def _load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            from . import build
            build.build_gen()
        _lib = C.CDLL(LIB_PATH)
        _lib.b200gen_generate.argtypes = [C.c_int, C.c_uint64, C.c_uint64, C.c_int, C.c_void_p,
                                          C.c_uint64, C.POINTER(C.c_uint64)]
        _lib.b200gen_generate.restype = C.c_uint64
    return _lib


def generate(kind, seed, target_bytes, ncols=128, out=None):
    """Returns (numpy uint8 array of text, rows). Text is whole lines and ends with a newline.
    `out` may be a preallocated uint8 array (e.g. pinned) with slack >= 2 MiB beyond target."""
    lib = _load()
    kind = KINDS.get(kind, kind)
    slack = 4 << 20
    if out is None:
        out = np.empty(int(target_bytes) + slack, dtype=np.uint8)
    rows = C.c_uint64(0)
    n = lib.b200gen_generate(kind, seed, int(target_bytes), ncols, out.ctypes.data, out.nbytes,
                             C.byref(rows))
    return out[:n], int(rows.value)
