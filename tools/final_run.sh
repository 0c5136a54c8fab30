# round-end check on one GPU: parity suite, both bench arms, per-kernel times of the five shapes
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu > gpurun_out/final_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/final_tests.log
timeout 600 python tools/first_pass.py > gpurun_out/first_pass.log 2>&1; echo "first_pass rc=$?" >> gpurun_out/final_tests.log
