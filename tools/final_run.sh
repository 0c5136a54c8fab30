# round-end evidence on one GPU: both bench arms and the ncu captures of the final binary
# (tools/refresh_profiles.sh turns the result into profiles/)
mkdir -p gpurun_out
timeout 900 python bench.py --impl reference > gpurun_out/final_bench_reference.json 2> gpurun_out/final_bench_reference.err; echo "ref rc=$?" > gpurun_out/final_tests.log
timeout 900 python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; echo "bench rc=$?" >> gpurun_out/final_tests.log
timeout 1500 bash tools/run_profiles.sh > gpurun_out/final_profiles.log 2>&1; echo "profiles rc=$?" >> gpurun_out/final_tests.log
