#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_4_configs.py -q -x --timeout=300 2>&1 | tail -n 3
timeout 300 python bench.py --skip-cpu --workload c3 > gpurun_out/bench_c3.log 2>&1; tail -n 1 gpurun_out/bench_c3.log | grep -o '"value": [0-9.]*\|"ms_per_step": [0-9.]*\|"frac": [0-9.]*' | tr '\n' ' '; echo
