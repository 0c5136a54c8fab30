timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "golden or edge or fuzz or corpus_device or hand_over or tile_boundaries or timestamp" 2>&1 | tail -2
timeout 300 python tools/bench_formats.py --mb 500 --no-e2e 2>&1 | python tools/fmt_lines.py
