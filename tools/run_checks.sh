timeout 900 python bench.py > gpurun_out/bench_r02c.json 2> gpurun_out/bench_r02c.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_r02c.err
timeout 600 python bench.py --impl reference > gpurun_out/bench_r02c_ref.json 2> gpurun_out/bench_r02c_ref.err; echo "ref rc=$?"
timeout 300 python tools/bench_formats.py --mb 1000 > gpurun_out/r02_formats.jsonl 2>gpurun_out/r02_formats.err; python tools/fmt_lines.py < gpurun_out/r02_formats.jsonl
