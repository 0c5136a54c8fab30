# Turn what tools/final_run.sh brought back in gpurun_out/ into the tracked files under profiles/
# (run here, after the GPU call): ncu summaries, per-line tables, launch list, SASS listings of the
# in-tree library, bench lines.
set -e
cd "$(dirname "$0")/.."
python profiles/summarize_ncu.py gpurun_out/r02_split_libfm.ncu-rep profiles/r02_split_libfm 1000097056 > /dev/null
python profiles/summarize_ncu.py gpurun_out/r02_split_libsvm.ncu-rep profiles/r02_split_libsvm 1000203280 > /dev/null
python profiles/summarize_ncu.py gpurun_out/r02_split_csv.ncu-rep profiles/r02_split_csv 1000080464 > /dev/null
python profiles/summarize_ncu.py gpurun_out/r02_tiles_csv.ncu-rep profiles/r02_tiles_csv 1000080464 > /dev/null
for n in r02_split_libfm r02_split_libsvm r02_split_csv r02_tiles_csv; do
  python profiles/lines_by_source.py gpurun_out/$n.ncu-rep 45 > profiles/$n.lines.txt
  python profiles/stalls_by_source.py gpurun_out/$n.ncu-rep >> profiles/$n.lines.txt 2> /dev/null || true
done
python - <<'EOF'
import json
out = []
for n in ("r02_split_libfm", "r02_split_libsvm", "r02_split_csv", "r02_tiles_csv"):
    d = json.load(open("profiles/%s.metrics.json" % n))
    seen = set()
    for k in (d if isinstance(d, list) else d.get("kernels", [d])):
        if k["kernel"] in seen:
            continue
        seen.add(k["kernel"])
        k.setdefault("capture", n)
        out.append(k)
json.dump(out, open("profiles/current_ncu.metrics.json", "w"), indent=1)
for k in out:
    print("%-62s %.3f ms  issue %.1f %%  ALU %.1f %%  %.3f warp instr / byte  %.2f GB DRAM" % (
        k["kernel"][:62], k["gpu__time_duration.sum"] * 1e3,
        k["smsp__issue_active.avg.pct_of_peak_sustained_active"],
        k["sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active"],
        k["smsp__inst_executed.sum"] / k["text_bytes"], k["dram_bytes_per_launch"] / 1e9))
EOF
cp gpurun_out/r02_launches.csv profiles/r02_launches.csv
cp gpurun_out/final_bench.json profiles/r02_bench.json
cp gpurun_out/final_bench_reference.json profiles/r02_bench_reference.json
gen() {
  cuobjdump -sass -fun "$1" dmlc_core_b200/lib/libdmlc_b200.so 2> /dev/null | grep -v "^\s*/\* 0x" |
    sed 's#/\* 0x[0-9a-f]* \*/##' | sed -n '/Function :/,$p' |
    sed '/^Fatbin\|^====\|^arch =\|^code version\|^host =\|^compile_size\|^identifier\|^$/d' > profiles/$2
}
gen _ZN5b200p13CollectRangesILi1EmLb0EEEvNS_11SplitParamsE r02_sass_collect_libfm_u64.txt
gen _ZN5b200p13ConvertRangesILi1EmEEvNS_11SplitParamsE r02_sass_convert_libfm_u64.txt
gen _ZN5b200p18CollectFieldRangesIjEEvNS_11SplitParamsE r02_sass_collect_csv_u32.txt
gen _ZN5b200p18ConvertFieldRangesIjEEvNS_11SplitParamsE r02_sass_convert_csv_u32.txt
gen _ZN5b200p10ParseTilesILi2EjLb0EEEvNS_10TileParamsE r02_sass_parsetiles_csv_u32.txt
