timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "golden or edge or fuzz or corpus_device" 2>&1 | tail -3
echo "== W=8 sub=2048"; timeout 300 python tools/bench_formats.py --mb 500 --no-e2e 2>&1 | python tools/fmt_lines.py
echo "== W=16 sub=2048"; DMLC_B200_LIB=$PWD/dmlc_core_b200/lib/libdmlc_b200_w16.so timeout 300 python tools/bench_formats.py --mb 500 --no-e2e 2>&1 | python tools/fmt_lines.py
echo "== W=16 parity"; DMLC_B200_LIB=$PWD/dmlc_core_b200/lib/libdmlc_b200_w16.so timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "golden or edge or corpus_device" 2>&1 | tail -3
