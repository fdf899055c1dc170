#!/bin/bash
mkdir -p gpurun_out
for o in 0 1 0 1 0 1; do for m in 31 0; do
  VQ_SPLIT_OVERLAP=$o VQ_PDL=$m timeout 300 python bench.py --skip-cpu --workload c3 --steps 100 > gpurun_out/b.log 2>&1; echo "c3 overlap=$o pdl=$m $(tail -n 1 gpurun_out/b.log | grep -o '"ms_per_step": [0-9.]*' | head -1)" | tee -a gpurun_out/exp2.log
done; done
