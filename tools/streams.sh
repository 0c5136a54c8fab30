# device-resident rate of bench.py at 1..4 handles/streams
mkdir -p gpurun_out; : > gpurun_out/streams.log
for k in 1 2 3 4; do
  timeout 300 python bench.py --no-e2e --no-cpu-baseline --no-shapes --streams $k 2>> gpurun_out/streams.err | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print($k, round(d['value'], 1), round(d['ms_per_step'], 3), round(d['roofline']['whole_step']['frac'], 4))" >> gpurun_out/streams.log
done
