# per-kernel times of the five shapes at several range sizes (1000 MB and 64 MB windows)
mkdir -p gpurun_out; : > gpurun_out/ranges.log
for mb in 1000 64; do for r in 4096 8192 16384 32768; do
  echo "== $mb MB, range $r" >> gpurun_out/ranges.log
  timeout 300 python tools/bench_formats.py --mb $mb --no-e2e --reps 7 --range-bytes $r 2>> gpurun_out/ranges.err | python tools/fmt_lines.py >> gpurun_out/ranges.log
done; done
