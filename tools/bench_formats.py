#!/usr/bin/env python
"""Device-resident throughput of the five BASELINE.json corpus shapes (one B200), per kernel.
Not the contract bench (that is bench.py on configs[1]); this is the per-format evidence for
profiles/.   python tools/bench_formats.py [--mb 512] [--range-bytes 0]"""
import argparse
import json
import os
import sys

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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, default=512)
    ap.add_argument("--range-bytes", type=int, default=0)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--only", default="", help="substring of the shape name to run")
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
        print(json.dumps(out[-1]))
        p.close()
        del dev
    return out


if __name__ == "__main__":
    main()
