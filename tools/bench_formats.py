#!/usr/bin/env python
"""Throughput of the five BASELINE.json corpus shapes on one B200: device-resident per kernel, and
end to end through b200p_submit / b200p_next (pinned host text in, pinned host CSR out).
Not the contract bench (that is bench.py on configs[1]); this is the per-format evidence for
profiles/.   python tools/bench_formats.py [--mb 512] [--range-bytes 0] [--no-e2e]"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
from dmlc_core_b200 import Parser, datagen  # noqa: E402

SHAPES = [  # BASELINE config, generator kind, format, index bytes, label column
    ("cfg1 libsvm u32", "libsvm", "libsvm", 4, -1),
    ("cfg2 csv u32", "csv", "csv", 4, 0),
    ("cfg3 libfm u64", "libfm", "libfm", 8, -1),
    ("cfg4 libsvm weight+qid u32", "libsvm_qid", "libsvm", 4, -1),
    ("cfg5 criteo libsvm u64", "criteo", "libsvm", 8, -1),
]


def end_to_end(data, fmt, ib, lc, args):
    """The same text from pinned host memory through b200p_submit / b200p_next, `passes` times."""
    host = torch.from_numpy(data).pin_memory()
    raw = host.numpy()
    base = host.data_ptr()
    step = args.chunk_mb << 20
    cuts, pos = [0], 0
    while pos < raw.size:
        pos = min(raw.size, pos + step)
        while pos < raw.size and raw[pos - 1] != 10:
            pos -= 1
        cuts.append(pos)
    depth = 3
    p = Parser(fmt, index_bytes=ib, label_column=lc, chunk_bytes=step, depth=depth)

    def one_pass():
        rows = 0
        for a, b in zip(cuts[:-1], cuts[1:]):
            while p.pending() >= depth:
                rows += p.next_raw().rows
            p.submit_in_place(base + a, b - a)
        while p.pending():
            rows += p.next_raw().rows
        return rows

    one_pass()
    up0, down0 = p.copied_bytes()
    t0 = time.perf_counter()
    passes = 4
    for _ in range(passes):
        one_pass()
    dt = (time.perf_counter() - t0) / passes
    up1, down1 = p.copied_bytes()
    p.close()
    return {"e2e_GBps": raw.size / dt / 1e9, "e2e_h2d_bytes": (up1 - up0) // passes,
            "e2e_d2h_bytes": (down1 - down0) // passes}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, default=512)
    ap.add_argument("--range-bytes", type=int, default=0)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--only", default="", help="substring of the shape name to run")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--chunk-mb", type=int, default=64)
    args = ap.parse_args()
    peak = 6540.8
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    out = []
    for name, kind, fmt, ib, lc in SHAPES:
        if args.only and args.only not in name:
            continue
        data, rows = datagen.generate(kind, 1, args.mb * 1000 * 1000)
        n = data.size
        pad = (-n) % 16
        buf = np.concatenate([data, np.full(pad, 10, dtype=np.uint8)])
        dev = torch.from_numpy(buf).cuda()
        stream = torch.cuda.current_stream()
        p = Parser(fmt, index_bytes=ib, label_column=lc, range_bytes=args.range_bytes,
                   stream=stream.cuda_stream)
        blk = None
        for _ in range(3):
            blk = p.parse_device(dev.data_ptr(), dev.numel())
        p.set_profiling(True)
        for _ in range(args.reps):
            blk = p.parse_device(dev.data_ptr(), dev.numel())
        torch.cuda.synchronize()
        ms = [p.kernel_time(k)[0] / max(p.kernel_time(k)[1], 1) for k in range(3)]
        p.set_profiling(False)
        b_alg = dev.numel() + 8 * (blk.rows + 1) + 4 * blk.rows + 4 * blk.n_weight + 8 * blk.n_qid + \
            blk.nnz * ib + blk.n_field * ib + 4 * blk.n_value
        total = sum(ms)
        out.append({"shape": name, "text_bytes": int(dev.numel()), "rows": int(blk.rows),
                    "nnz": int(blk.nnz), "count_ms": ms[0], "scan_ms": ms[1], "emit_ms": ms[2],
                    "input_GBps": dev.numel() / total / 1e6, "b_alg": int(b_alg),
                    "emit_frac_of_hbm_peak": b_alg / ms[2] / 1e6 / peak,
                    "all_kernels_frac_of_hbm_peak": b_alg / total / 1e6 / peak})
        p.close()
        del dev
        if not args.no_e2e:
            out[-1].update(end_to_end(data, fmt, ib, lc, args))
        print(json.dumps(out[-1]))
    return out


if __name__ == "__main__":
    main()
