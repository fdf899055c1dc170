#!/bin/bash
# quick correctness + kernel timing trip
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x --timeout=300 > gpurun_out/tests.log 2>&1; echo "tests exit $?" >> gpurun_out/tests.log
tail -n 4 gpurun_out/tests.log
for v in "VQ_ASSIGN_PAIR=1" "VQ_ASSIGN_PAIR=0" "VQ_ASSIGN_EPI_WARPS=16"; do
  echo "$v" | tee -a gpurun_out/kern.log
  env $v timeout 300 python tools/bench_kernels.py c2 c4 c1 2>&1 | tee -a gpurun_out/kern.log
done
