# per-kernel times of the five shapes for every library under dmlc_core_b200/lib/variants/ and for the in-tree one
mkdir -p gpurun_out; : > gpurun_out/variants.log
for lib in dmlc_core_b200/lib/libdmlc_b200.so dmlc_core_b200/lib/variants/*.so; do
  echo "== $lib" >> gpurun_out/variants.log
  DMLC_B200_LIB=$PWD/$lib timeout 300 python tools/bench_formats.py --mb 1000 --no-e2e --reps 7 2>> gpurun_out/variants.err | python tools/fmt_lines.py >> gpurun_out/variants.log
done
