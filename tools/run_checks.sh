timeout 1700 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
timeout 900 python bench.py > gpurun_out/bench_r02b.json 2> gpurun_out/bench_r02b.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_r02b.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_r02b_ref.json 2> gpurun_out/bench_r02b_ref.err; echo "ref rc=$?"
