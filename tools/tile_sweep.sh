
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "golden or edge or fuzz or corpus_device" > gpurun_out/pt2.log 2>&1; tail -5 gpurun_out/pt2.log
for cfg in "2048 4096" "2048 2048" "3072 4096" "4096 4096" "1024 2048"; do set -- $cfg; echo "== sub=$1 over=$2"; B200P_SUB_BYTES=$1 B200P_OVER_BYTES=$2 timeout 300 python tools/bench_formats.py --mb 500 --no-e2e 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except: print(l.strip()[:200]); continue
    print('%-28s emit %.3f ms  %.0f GB/s frac %.3f'%(d['shape'],d['emit_ms'],d['input_GBps'],d['all_kernels_frac_of_hbm_peak']))
"; done
