"""one line per shape from the JSON lines of tools/bench_formats.py (stdin)"""
import json
import sys
for line in sys.stdin:
    try:
        d = json.loads(line)
    except ValueError:
        print(line.strip()[:200])
        continue
    print("%-28s count %.3f emit %.3f ms  %.0f GB/s  frac %.3f" % (
        d["shape"], d["count_ms"], d["emit_ms"], d["input_GBps"], d["all_kernels_frac_of_hbm_peak"]))
