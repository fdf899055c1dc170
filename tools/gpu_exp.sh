#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x --timeout=300 2>&1 | tail -n 3
timeout 300 python tools/bench_kernels.py c4 c5 c2 2>&1 | tail -n 3 | cut -c1-170
timeout 300 python bench.py --skip-cpu --workload c4 2>&1 | tail -n 1 | grep -o '"value": [0-9.]*\|"ms_per_step": [0-9.]*\|"frac": [0-9.]*' | tr '\n' ' '; echo
