#!/usr/bin/env python
"""Parser::Create from a file in /dev/shm at several reader-thread counts / chunk sizes / depths
(the knobs of the host feeder): GB/s per setting, one JSON line each."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

if len(sys.argv) > 1 and sys.argv[1] == "child":
    from dmlc_core_b200 import hostapi
    uri, fmt, ib = sys.argv[2], sys.argv[3], int(sys.argv[4])
    first = hostapi.time_parser(uri, 0, 1, fmt, ib, passes=1)        # first pass of the process: allocations, module load, mapping
    got = hostapi.time_parser(uri, 0, 1, fmt, ib, passes=3)
    print(json.dumps({"GBps": got["bytes_read"] / got["seconds"] / 1e9,
                      "first_pass_GBps": first["bytes_read"] / (first["seconds"] + first["create_seconds"]) / 1e9}))
    sys.exit(0)

from dmlc_core_b200 import datagen  # noqa: E402
kind, fmt, ib = "libfm", "libfm", 8
data, rows = datagen.generate(kind, 1, 1000 * 1000 * 1000)
path = "/dev/shm/b200_sweep.txt"
with open(path, "wb") as f:
    for _ in range(4):
        data.tofile(f)
for threads, chunk, depth, use_map in ((12, 64, 3, 0), (12, 64, 3, 1), (12, 64, 3, 0), (12, 64, 3, 1), (10, 64, 3, 1),
                                       (12, 64, 4, 1), (12, 128, 3, 1)):
    env = dict(os.environ, DMLC_B200_READ_THREADS=str(threads), DMLC_B200_CHUNK_MB=str(chunk),
               DMLC_B200_DEPTH=str(depth), DMLC_B200_READ_MMAP=str(use_map))
    r = subprocess.run([sys.executable, __file__, "child", path, fmt, str(ib)], env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    line = r.stdout.strip().splitlines()[-1] if r.stdout.strip() else r.stderr[-300:]
    print(json.dumps({"read_threads": threads, "chunk_mb": chunk, "depth": depth, "mmap": use_map, "result": line}))
os.remove(path)
