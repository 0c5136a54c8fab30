# round-end check on one GPU: parity suite and per-kernel times of the five shapes
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/final_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/final_tests.log
timeout 600 python tools/bench_formats.py --mb 1000 > gpurun_out/final_formats.jsonl 2> gpurun_out/final_formats.err; echo "formats rc=$?" >> gpurun_out/final_tests.log
