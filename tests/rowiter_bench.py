"""RowBlockIter<uint32>::Create(uri) load time (the XGBoost-style in-memory load)
through this package's C++ layer and through the compiled reference, same file.
Run on the GPU box:  python tests/rowiter_bench.py [copies]"""
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import ctypes as C

from dmlc_core_b200 import datagen, hostapi
from oracle.pyoracle import Reference


def main():
    copies = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    for kind, type_, uri_tail in (("csv", "csv", "?format=csv&label_column=0"), ("libsvm", "libsvm", "")):
        path = "/dev/shm/b200_iter.%s" % kind
        data, rows = datagen.generate(kind, 7, 512 << 20)
        with open(path, "wb") as f:
            for _ in range(copies):
                f.write(data.tobytes())
        size = os.path.getsize(path)
        lib = hostapi.load()
        for rep in range(2):
            b, ncol = hostapi.Block(), C.c_uint64(0)
            err = C.create_string_buffer(2048)
            t0 = time.perf_counter()
            h = lib.dmlcb_rowiter_create((path + uri_tail).encode(), 0, 1, type_.encode(), 4,
                                         C.byref(b), C.byref(ncol), err, 2048)
            dt = time.perf_counter() - t0
            assert h, err.value
            lib.dmlcb_rowiter_free(h)
            print("%s b200 RowBlockIter::Create  %.2f GB/s  (%.2f s, NumCol %d)" %
                  (kind, size / dt / 1e9, dt, ncol.value), flush=True)
        ref = Reference.load()
        if ref is not None:
            t0 = time.perf_counter()
            out, nc = ref.iter_uri(path + uri_tail, 0, 1, type_, 4)
            dt = time.perf_counter() - t0
            print("%s reference RowBlockIter::Create + copy out  %.2f GB/s (%.2f s, NumCol %d)" %
                  (kind, size / dt / 1e9, dt, nc), flush=True)
        os.remove(path)


if __name__ == "__main__":
    main()
