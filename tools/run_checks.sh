# GPU box: the new tests, the probe of the host read path, and a first bench line
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_cpp_parser.py tests/test_zz_no_alloc.py -x -q -m gpu -k "hands_over or tile_boundaries or device_csr_out or consumer_is_busy or rewind or noalloc" 2>&1 | tail -8
timeout 300 python tools/probe_mmap_register.py 2000 > gpurun_out/probe_mmap.json 2> gpurun_out/probe_mmap.err; tail -c 1500 gpurun_out/probe_mmap.json; tail -3 gpurun_out/probe_mmap.err
timeout 900 python bench.py > gpurun_out/bench_r02a.json 2> gpurun_out/bench_r02a.err; echo "bench rc=$?"; tail -5 gpurun_out/bench_r02a.err; head -c 3000 gpurun_out/bench_r02a.json
