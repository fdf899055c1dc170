#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x --timeout=300 2>&1 | tail -n 1
for i in 1 2 3 4 5; do
  timeout 300 python bench.py --skip-cpu --workload c3 > gpurun_out/bench_c3.log 2>&1; echo "c3 $(tail -n 1 gpurun_out/bench_c3.log | grep -o '"ms_per_step": [0-9.]*' | head -1)" | tee -a gpurun_out/exp2.log
done
timeout 300 python bench.py --skip-cpu > gpurun_out/b.log 2>&1; echo "c2 $(tail -n 1 gpurun_out/b.log | grep -o '"ms_per_step": [0-9.]*' | head -1)" | tee -a gpurun_out/exp2.log
