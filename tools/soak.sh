# long fuzz run of the parity suite (GPU): B200_FUZZ_SEED x B200_FUZZ_CASES on the default and the one-launch path
for seed in 11 12 13; do
  B200_FUZZ_SEED=$seed B200_FUZZ_CASES=4000 timeout 1500 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "fuzz and (default or tiles)" 2>&1 | tail -2
done
python __graft_entry__.py smoke 2>&1 | tail -3
