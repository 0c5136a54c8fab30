"""GPU test of b200p_parse_device_noalloc, kept in a file pytest collects last so that a run
stopping at the first failure has already been through every other test.

The entry point exists for hosts that may not allocate GPU memory while they run -- a process
whose whole card belongs to a pool allocator (RMM under XGBoost, a framework's caching allocator)
or one busy with NCCL collectives, where cudaMalloc/cudaFree stalls or fails.  Here torch plays
that pool: the test allocates the text and every CSR array up front, the call itself only uses
the library's fixed scratch.  Its output must equal b200p_parse_device's bit for bit (the same
tolerance as every parity test in this suite: 0 ulp), on two calls in a row, because the second
call finds the first call's counters still sitting in the shared scratch.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

NAMES = ("offset", "label", "weight", "qid", "field", "index", "value")


def _pool(torch, rows, nnz, ib):
    idt = torch.int64 if ib == 8 else torch.int32
    return {
        "offset": torch.full((rows + 1,), -1, dtype=torch.int64, device="cuda"),
        "label": torch.full((rows,), -1.0, dtype=torch.float32, device="cuda"),
        "weight": torch.full((rows,), -1.0, dtype=torch.float32, device="cuda"),
        "qid": torch.full((rows,), -1, dtype=torch.int64, device="cuda"),
        "field": torch.full((nnz,), -1, dtype=idt, device="cuda"),
        "index": torch.full((nnz,), -1, dtype=idt, device="cuda"),
        "value": torch.full((nnz,), -1.0, dtype=torch.float32, device="cuda"),
    }


def _request(capi, torch, fmt, ib, lc, text, pool, rows, nnz):
    rq = capi.NoAllocRequest()
    rq.format = capi.FORMATS[fmt]
    rq.index_bytes = ib
    rq.label_column = lc
    rq.device = torch.cuda.current_device()
    rq.stream = torch.cuda.current_stream().cuda_stream
    rq.d_text = text.data_ptr()
    rq.nbytes = text.numel()
    for k in NAMES:
        setattr(rq, "d_" + k, pool[k].data_ptr())
    rq.cap_rows = rows
    rq.cap_nnz = nnz
    return rq


def _as_numpy(blk, pool, ib):
    idt = np.uint64 if ib == 8 else np.uint32
    n = {"offset": blk.rows + 1, "label": blk.rows, "weight": blk.n_weight, "qid": blk.n_qid,
         "field": blk.n_field, "index": blk.nnz, "value": blk.n_value}
    view = {"offset": np.uint64, "label": np.float32, "weight": np.float32, "qid": np.uint64,
            "field": idt, "index": idt, "value": np.float32}
    out = {k: pool[k][:int(n[k])].cpu().numpy().view(view[k]) for k in NAMES}
    out["error"] = 0
    return out


@pytest.mark.parametrize("kind,fmt,ib,lc", [("libsvm_qid", "libsvm", 4, -1), ("libfm", "libfm", 8, -1),
                                            ("csv", "csv", 4, 0)])
def test_noalloc_matches_parse_device_twice(cuda_lib, oracle, kind, fmt, ib, lc):
    import torch
    from dmlc_core_b200 import Parser, capi, datagen
    from gpu_util import parse_on_device
    from oracle.pyoracle import same_rowblocks

    # two different chunks: the second call must not be disturbed by what the first one left
    texts = [datagen.generate(kind, 31, 3 << 20)[0], datagen.generate(kind, 32, 1 << 20)[0]]
    ref_parser = Parser(fmt, index_bytes=ib, label_column=lc)
    expected = [parse_on_device(ref_parser, t) for t in texts]
    ref_parser.close()
    # the existing entry point is itself pinned to the CPU oracle on these very chunks
    for t, e in zip(texts, expected):
        assert same_rowblocks(e, oracle.parse_buffer(fmt, t.tobytes(), ib, lc)) is None

    rows = max(len(e["label"]) for e in expected)
    nnz = max(len(e["index"]) for e in expected)
    pool = _pool(torch, rows, nnz, ib)            # every allocation happens here, before the calls
    dev_texts = [torch.from_numpy(t.copy()).cuda() for t in texts]
    torch.cuda.synchronize()
    for call, (text, exp) in enumerate(zip(dev_texts, expected)):
        before = torch.cuda.memory_allocated()
        free_before = torch.cuda.mem_get_info()[0]
        rc, blk, msg = capi.parse_device_noalloc(_request(capi, torch, fmt, ib, lc, text, pool, rows, nnz))
        assert rc == capi.OK, msg
        assert torch.cuda.memory_allocated() == before
        if call == 1:   # kernels and module are loaded by now: the device's free memory is untouched
            assert torch.cuda.mem_get_info()[0] >= free_before
        assert same_rowblocks(_as_numpy(blk, pool, ib), exp) is None


def test_noalloc_limits_are_reported(cuda_lib):
    import torch
    from dmlc_core_b200 import capi
    text = torch.zeros(64, dtype=torch.uint8, device="cuda")
    pool = _pool(torch, 4, 4, 4)
    rq = _request(capi, torch, "libsvm", 4, -1, text, pool, 4, 4)
    rq.nbytes = capi.NOALLOC_MAX_BYTES + 1      # refused before anything touches the device
    rc, _, msg = capi.parse_device_noalloc(rq)
    assert rc == capi.E_INVALID and "exceeds" in msg
    # arrays too small: nothing written, needed sizes reported
    data = np.frombuffer(b"1 1:1 2:2 3:3\n0 4:4 5:5 6:6\n" * 8, dtype=np.uint8)
    t = torch.from_numpy(data.copy()).cuda()
    rc, blk, msg = capi.parse_device_noalloc(_request(capi, torch, "libsvm", 4, -1, t, pool, 4, 4))
    assert rc == capi.E_CAPACITY and blk.rows == 16 and blk.nnz == 48, msg
    assert int(pool["index"][0].item()) == -1
    # the same for CSV, whose labels are cleared by a kernel of their own before pass 2
    data = np.frombuffer(b"1,2,3\n4,5,6\n" * 8, dtype=np.uint8)
    t = torch.from_numpy(data.copy()).cuda()
    rc, blk, msg = capi.parse_device_noalloc(_request(capi, torch, "csv", 4, -1, t, pool, 4, 4))
    assert rc == capi.E_CAPACITY and blk.rows == 16 and blk.nnz == 48, msg
    assert float(pool["label"][0].item()) == -1.0 and float(pool["value"][0].item()) == -1.0
