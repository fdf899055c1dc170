#!/bin/bash
# experiment trip: parity tests, host-side profile of the small-batch call, short bench lines
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x --timeout=300 2>&1 | tail -n 2
timeout 300 python tools/host_profile.py 2>&1 | cut -c1-150 | head -24 | tee gpurun_out/host_profile.log
for w in c1 c1 c2; do
  timeout 300 python bench.py --skip-cpu --workload $w --steps 200 > gpurun_out/bench_$w.log 2>&1; echo "$w $(tail -n 1 gpurun_out/bench_$w.log | grep -o '"ms_per_step": [0-9.]*' | tr '\n' ' ')"
done
