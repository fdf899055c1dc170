#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x --timeout=300 2>&1 | tail -n 1
for w in c2 c4; do for d in gaussian trained; do
  timeout 300 python bench.py --skip-cpu --workload $w --data $d > gpurun_out/b.log 2>&1; echo "$w $d $(tail -n 1 gpurun_out/b.log | grep -o '"ms_per_step": [0-9.]*\|"ms_per_launch": [0-9.]*' | tr '\n' ' ')" | tee -a gpurun_out/exp2.log
done; done
