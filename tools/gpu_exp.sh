#!/bin/bash
mkdir -p gpurun_out
for m in 3 1; do
  echo "mode=$m" | tee -a gpurun_out/exp.log
  VQ_ASSIGN_MODE=$m timeout 300 python tools/bench_kernels.py c4 2>&1 | cut -c1-70 | tee -a gpurun_out/exp.log
done
python bench.py --skip-cpu 2>&1 | tail -n 1 | cut -c1-330
