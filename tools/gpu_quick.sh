#!/bin/bash
# quick correctness + kernel timing trip
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x --timeout=300 > gpurun_out/tests.log 2>&1; echo "tests exit $?" >> gpurun_out/tests.log
tail -n 4 gpurun_out/tests.log
for v in "16 4096" "16 1" "16 1000000" "8 4096"; do
  set -- $v
  echo "epi_warps=$1 branch_min_k=$2" | tee -a gpurun_out/kern.log
  VQ_ASSIGN_EPI_WARPS=$1 VQ_ASSIGN_BRANCH_MIN_K=$2 timeout 300 python tools/bench_kernels.py c2 c4 c1 2>&1 | tee -a gpurun_out/kern.log
done
