"""Helpers for the GPU tests: everything calls the product through the C ABI (capi.Parser)."""
import numpy as np
import torch


class _DevMem(object):
    def __init__(self, ptr, n, typestr):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": typestr,
                                         "data": (int(ptr), False), "version": 2}


def as_torch(ptr, n, np_dtype):
    """Zero-copy torch view of device memory owned by the parser handle."""
    if not ptr or n == 0:
        return torch.zeros(0, device="cuda")
    tmap = {np.dtype(np.float32): ("<f4", torch.float32), np.dtype(np.uint32): ("<i4", torch.int32),
            np.dtype(np.uint64): ("<i8", torch.int64)}
    ts, _ = tmap[np.dtype(np_dtype)]
    return torch.as_tensor(_DevMem(ptr, n, ts), device="cuda")


def to_device(data):
    """bytes -> CUDA uint8 tensor (torch allocations are >= 256-byte aligned)."""
    arr = np.frombuffer(bytes(data), dtype=np.uint8) if not isinstance(data, np.ndarray) else data
    if arr.size == 0:
        return torch.zeros(16, dtype=torch.uint8, device="cuda"), 0
    return torch.from_numpy(arr.copy()).cuda(), int(arr.size)


def parse_on_device(parser, data):
    t, n = to_device(data)
    blk = parser.parse_device(t.data_ptr(), n)
    out = parser.fetch(blk)
    del t
    return out
