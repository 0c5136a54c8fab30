#!/usr/bin/env python
"""Where the first pass of a fresh process goes: Parser::Create + one drain of a file in /dev/shm,
split into library load, CUDA start-up (first handle), the first pass of a new parser, the first
pass of a second parser in the warm process, and the steady state.  One JSON line per file size."""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

if len(sys.argv) > 1 and sys.argv[1] == "child":
    uri, fmt, ib = sys.argv[2], sys.argv[3], int(sys.argv[4])
    t0 = time.perf_counter()
    import numpy as np
    from dmlc_core_b200 import capi, hostapi
    capi.load()
    hostapi.load()
    t1 = time.perf_counter()
    tiny = capi.Parser(fmt, index_bytes=ib)
    tiny.parse_host(np.frombuffer(b"1 1:1:1\n", dtype=np.uint8))
    tiny.close()
    t2 = time.perf_counter()
    first = hostapi.time_parser(uri, 0, 1, fmt, ib, passes=1)
    t3 = time.perf_counter()
    second = hostapi.time_parser(uri, 0, 1, fmt, ib, passes=1)
    steady = hostapi.time_parser(uri, 0, 1, fmt, ib, passes=3)
    size = first["bytes_read"]
    print(json.dumps({
        "file_MB": round(size / 1e6),
        "load_libraries_ms": round((t1 - t0) * 1e3, 1),
        "cuda_startup_and_first_handle_ms": round((t2 - t1) * 1e3, 1),
        "first_parser_create_ms": round(first["create_seconds"] * 1e3, 1),
        "first_parser_pass_ms": round(first["seconds"] * 1e3, 1),
        "first_parser_wall_ms": round((t3 - t2) * 1e3, 1),
        "second_parser_create_ms": round(second["create_seconds"] * 1e3, 1),
        "second_parser_pass_ms": round(second["seconds"] * 1e3, 1),
        "steady_pass_ms": round(steady["seconds"] / 3 * 1e3, 1),
        "first_pass_GBps_of_a_warm_process": round(size / (t3 - t2) / 1e9, 2),
        "steady_GBps": round(size / (steady["seconds"] / 3) / 1e9, 2)}))
    sys.exit(0)

from dmlc_core_b200 import datagen  # noqa: E402
kind, fmt, ib = "libfm", "libfm", 8
data, rows = datagen.generate(kind, 1, 250 * 1000 * 1000)
path = "/dev/shm/b200_first_pass.txt"
for copies in (1, 4, 16):
    with open(path, "wb") as f:
        for _ in range(copies):
            data.tofile(f)
    env = dict(os.environ, B200P_TRACE="1")
    r = subprocess.run([sys.executable, __file__, "child", path, fmt, str(ib)], env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    print(r.stdout.strip().splitlines()[-1] if r.stdout.strip() else json.dumps({"error": r.stderr[-400:]}))
    for line in r.stderr.splitlines():
        if "trace" in line:
            print("#", line)
os.remove(path)
