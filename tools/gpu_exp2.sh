#!/bin/bash
mkdir -p gpurun_out
for rep in 1 2; do
for v in 0 31 30 1 6 24 14; do
  for w in c4; do
  VQ_PDL=$v timeout 300 python bench.py --skip-cpu --workload $w --steps 100 > gpurun_out/b.log 2>&1; echo "PDL=$v $w $(tail -n 1 gpurun_out/b.log | grep -o '"ms_per_step": [0-9.]*' | head -1)" | tee -a gpurun_out/exp2.log
  done
done
done
