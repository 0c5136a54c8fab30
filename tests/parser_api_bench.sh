#!/bin/bash
# Throughput of the C++ drop-in API (dmlc::Parser<unsigned>::Create -> Next loop) from files in
# the page cache, next to the compiled reference when oracle/_ref travelled along.
# Run on the GPU box from the repo root:  bash tests/parser_api_bench.sh [csv_copies] [svm_copies]
set -e
cd "$(dirname "$0")/.."
python tools/mkdata.py csv /dev/shm/b200_a.csv 512 "${1:-8}"
python tools/mkdata.py libsvm /dev/shm/b200_a.svm 512 "${2:-4}"
echo "== dmlc_core_b200 (3 passes each; the first pays the pinned allocations)"
tests/_build/libsvm_parser_driver "/dev/shm/b200_a.csv?format=csv&label_column=0" 0 1 csv 3
tests/_build/libsvm_parser_driver /dev/shm/b200_a.svm 0 1 libsvm 3
if [ -x tests/_build/libsvm_parser_driver_ref ]; then
  echo "== reference (OpenMP, all cores)"
  tests/_build/libsvm_parser_driver_ref "/dev/shm/b200_a.csv?format=csv&label_column=0" 0 1 csv 1
  tests/_build/libsvm_parser_driver_ref /dev/shm/b200_a.svm 0 1 libsvm 1
fi
rm -f /dev/shm/b200_a.csv /dev/shm/b200_a.svm
