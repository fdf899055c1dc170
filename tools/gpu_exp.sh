#!/bin/bash
# experiment trip: parity tests, kernel timings, short bench lines
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x --timeout=300 2>&1 | tail -n 2
timeout 300 python tools/bench_kernels.py c2 c4 c1 2>&1 | grep "stand-alone" | tee gpurun_out/kern.log
for w in c1 c2 c4; do
  timeout 300 python bench.py --skip-cpu --workload $w --steps 100 > gpurun_out/bench_$w.log 2>&1; echo "$w $(tail -n 1 gpurun_out/bench_$w.log | grep -o '"ms_per_step": [0-9.]*' | tr '\n' ' ')"
done
