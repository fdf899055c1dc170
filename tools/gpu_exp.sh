#!/bin/bash
# experiment trip: parity tests, short bench lines
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x --timeout=300 2>&1 | tail -n 2
for w in c1 c1 c2; do
  timeout 300 python bench.py --skip-cpu --workload $w --steps 200 > gpurun_out/bench_$w.log 2>&1; echo "$w $(tail -n 1 gpurun_out/bench_$w.log | grep -o '"ms_per_step": [0-9.]*' | tr '\n' ' ')"
done
