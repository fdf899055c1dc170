#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_4_configs.py -q -x --timeout=300 2>&1 | tail -n 2
for w in c2 c4; do
timeout 300 python bench.py --skip-cpu --steps 30 --workload $w --data trained > gpurun_out/bench_${w}_trained.log 2>&1; echo "exit $?" >> gpurun_out/bench_${w}_trained.log
tail -n 2 gpurun_out/bench_${w}_trained.log | grep -o '"value": [0-9.]*\|"ms_per_step": [0-9.]*\|"frac": [0-9.]*\|exit [0-9]*' | tr '\n' ' '; echo
done
