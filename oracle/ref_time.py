#!/usr/bin/env python
"""TEST / BENCH INFRASTRUCTURE ONLY: time the unmodified reference (oracle/_ref) in a process of
its own, so that OMP_NUM_THREADS -- the only thing that sizes the reference's OpenMP team, which
runs on the ThreadedParser's own thread (reference src/data/text_parser.h:98, parser.h:73-82) --
can be set per measurement.  Prints one JSON line.

    OMP_NUM_THREADS=k python oracle/ref_time.py <uri> <part> <nparts> <type> <index_bytes> <seconds> [passes]
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.pyoracle import Reference  # noqa: E402


def main():
    uri, part, nparts, type_, ib, seconds = (sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), sys.argv[4],
                                             int(sys.argv[5]), float(sys.argv[6]))
    fixed = int(sys.argv[7]) if len(sys.argv) > 7 else 0
    ref = Reference.load()
    if ref is None:
        print(json.dumps({"error": "oracle/_ref/libdmlc_ref.so missing"}))
        return
    first = ref.time_uri(uri, part, nparts, type_, ib, nthread=0, passes=1)
    if first["seconds"] <= 0:
        print(json.dumps({"error": "the reference failed on %s" % uri}))
        return
    passes = fixed if fixed > 0 else max(1, int(seconds / first["seconds"]))
    res = ref.time_uri(uri, part, nparts, type_, ib, nthread=0, passes=passes)
    per = res["seconds"] / passes
    print(json.dumps({"GBps": first["bytes"] / per / 1e9, "rows_per_s": first["rows"] / per,
                      "bytes": first["bytes"], "rows": first["rows"], "nnz": first["nnz"], "passes": passes,
                      "seconds_per_pass": per, "threads": ref.max_threads(),
                      "omp_num_threads": os.environ.get("OMP_NUM_THREADS")}))


if __name__ == "__main__":
    main()
