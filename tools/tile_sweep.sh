timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "golden or edge or fuzz or corpus or hand_over or tile_boundaries" 2>&1 | tail -3
timeout 300 python tools/bench_formats.py --mb 500 --no-e2e 2>&1 | python tools/fmt_lines.py
