#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x --timeout=300 > gpurun_out/tests.log 2>&1; echo "tests exit $?" >> gpurun_out/tests.log
tail -n 6 gpurun_out/tests.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -n 1
for w in c2 c4 c1; do timeout 300 python bench.py --skip-cpu --workload $w 2>&1 | tail -n 1 | grep -o '"value": [0-9.]*\|"ms_per_step": [0-9.]*\|"gpu_launches": [0-9]*' | tr '\n' ' '; echo; done
