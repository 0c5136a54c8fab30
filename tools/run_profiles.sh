# r02 profile captures (one B200): launch list of the bench command, full captures of the kernels of
# the headline shape (LibFM u64), of LibSVM u32 and of CSV (both of its paths), each behind a plain
# run of the same command line
CMD="python bench.py --gb 2 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-shapes"
$CMD > gpurun_out/prof_plain.json 2> gpurun_out/prof_plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches.csv $CMD > gpurun_out/prof_l.log 2>&1
$CMD > gpurun_out/prof_plain.json 2> gpurun_out/prof_plain.err && \
ncu --set full --clock-control none --import-source on -k regex:Ranges -s 12 -c 2 -o gpurun_out/r02_split_libfm -f $CMD > gpurun_out/prof_a.log 2>&1
CMD2="$CMD --shape 1"
$CMD2 > gpurun_out/prof_plain2.json 2> gpurun_out/prof_plain2.err && \
ncu --set full --clock-control none --import-source on -k regex:Ranges -s 12 -c 2 -o gpurun_out/r02_split_csv -f $CMD2 > gpurun_out/prof_b.log 2>&1
B200P_SPLIT_CSV=0 $CMD2 > gpurun_out/prof_plain2t.json 2> gpurun_out/prof_plain2t.err && \
B200P_SPLIT_CSV=0 ncu --set full --clock-control none --import-source on -k regex:ParseTiles -s 6 -c 1 -o gpurun_out/r02_tiles_csv -f $CMD2 > gpurun_out/prof_bt.log 2>&1
CMD3="$CMD --shape 0"
$CMD3 > gpurun_out/prof_plain3.json 2> gpurun_out/prof_plain3.err && \
ncu --set full --clock-control none --import-source on -k regex:Ranges -s 12 -c 2 -o gpurun_out/r02_split_libsvm -f $CMD3 > gpurun_out/prof_c.log 2>&1
for f in a b bt c; do tail -n 1 gpurun_out/prof_$f.log; done
