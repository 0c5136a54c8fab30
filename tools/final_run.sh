# round-end check on one GPU: parity suite, per-kernel times, first-pass decomposition
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/final_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/final_tests.log
timeout 600 python tools/bench_formats.py --mb 1000 > gpurun_out/final_formats.jsonl 2> gpurun_out/final_formats.err; echo "formats rc=$?" >> gpurun_out/final_tests.log
timeout 600 python tools/first_pass.py > gpurun_out/first_pass.log 2>&1; echo "first_pass rc=$?" >> gpurun_out/final_tests.log
