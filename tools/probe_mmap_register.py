#!/usr/bin/env python
"""Probe for the host side of Parser::Create: can the H2D copies read the page cache in place?

  1. mmap(PROT_READ, MAP_SHARED) of a file in /dev/shm + cudaHostRegister(..., cudaHostRegisterReadOnly):
     does it succeed, and what does the registration cost per GB?
  2. the C ABI fed from that mapping (b200p_submit copies to the device straight from memory the
     caller pinned) against the same text fed from the slots' own pinned staging buffers.
  3. O_DIRECT reads of a file in /tmp into a pinned buffer (tmpfs does not take O_DIRECT).

Prints one JSON line; numbers go into DESIGN.md."""
import ctypes
import json
import mmap
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
from dmlc_core_b200 import Parser, datagen  # noqa: E402

GB = 1e9
out = {}
mb = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
data, rows = datagen.generate("libfm", 1, mb * 1000 * 1000)
n = data.size
path = "/dev/shm/b200_probe.txt"
data.tofile(path)
rt = torch.cuda.cudart()
attr_ro = None
try:
    # cudaDevAttrHostRegisterReadOnlySupported = 113
    attr_ro = torch.cuda.cudart().cudaDeviceGetAttribute(113, 0) if hasattr(rt, "cudaDeviceGetAttribute") else None
except Exception as e:   # noqa: BLE001
    attr_ro = "query failed: %s" % e
out["cudaDevAttrHostRegisterReadOnlySupported"] = attr_ro

fd = os.open(path, os.O_RDONLY)
mm = mmap.mmap(fd, n, flags=mmap.MAP_SHARED, prot=mmap.PROT_READ)
addr = ctypes.addressof(ctypes.c_char.from_buffer_copy(b"x"))  # placeholder
buf = np.frombuffer(mm, dtype=np.uint8)
addr = buf.ctypes.data
# touch the pages once so the mapping is populated (page cache is warm: the file was just written)
_ = int(buf[::4096].sum())
for name, flags in (("readonly", 0x08), ("readonly_portable", 0x09), ("default", 0x00)):
    t0 = time.perf_counter()
    rc = rt.cudaHostRegister(addr, n, flags)
    dt = time.perf_counter() - t0
    out["register_" + name] = {"rc": int(rc), "seconds": dt, "seconds_per_GB": dt / (n / GB)}
    if int(rc) == 0:
        registered = name
        break
else:
    registered = None
out["registered_with"] = registered


def cuts_of(arr, size, step):
    cuts, pos = [0], 0
    while pos < size:
        pos = min(size, pos + step)
        while pos < size and arr[pos - 1] != 10:
            pos -= 1
        cuts.append(pos)
    return cuts


def run(base_addr, arr, passes=3):
    cuts = cuts_of(arr, n, 64 << 20)
    p = Parser("libfm", index_bytes=8, chunk_bytes=64 << 20, depth=3)

    def one():
        got = 0
        for a, b in zip(cuts[:-1], cuts[1:]):
            while p.pending() >= 3:
                got += p.next_raw().rows
            p.submit_in_place(base_addr + a, b - a)
        while p.pending():
            got += p.next_raw().rows
        return got
    assert one() == rows
    t0 = time.perf_counter()
    for _ in range(passes):
        one()
    dt = (time.perf_counter() - t0) / passes
    p.close()
    return n / dt / GB


if registered is not None:
    out["cabi_from_registered_mmap_GBps"] = run(addr, buf)
    rt.cudaHostUnregister(addr)
host = torch.from_numpy(data.copy()).pin_memory()
out["cabi_from_pinned_copy_GBps"] = run(host.data_ptr(), host.numpy())
out["cabi_from_unregistered_mmap_GBps (staged by memcpy)"] = run(addr, buf, passes=1)

# O_DIRECT into a pinned buffer, file on /tmp
try:
    p2 = "/tmp/b200_probe_direct.txt"
    data[:512 * 1000 * 1000].tofile(p2)
    os.sync()
    size2 = os.path.getsize(p2) // 4096 * 4096
    fd2 = os.open(p2, os.O_RDONLY | os.O_DIRECT)
    pinned = torch.empty(64 << 20, dtype=torch.uint8).pin_memory()
    view = memoryview(pinned.numpy())
    t0 = time.perf_counter()
    off = 0
    while off < size2:
        want = min(64 << 20, size2 - off)
        got = os.preadv(fd2, [view[:want]], off)
        if got <= 0:
            break
        off += got
    dt = time.perf_counter() - t0
    os.close(fd2)
    out["o_direct_tmp_GBps_single_thread"] = off / dt / GB
    os.remove(p2)
except Exception as e:   # noqa: BLE001
    out["o_direct_tmp"] = "failed: %s" % e
os.remove(path)
print(json.dumps(out))
