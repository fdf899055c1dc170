#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x --timeout=300 > gpurun_out/tests.log 2>&1; echo "tests exit $?" >> gpurun_out/tests.log
tail -n 15 gpurun_out/tests.log
timeout 300 python bench.py --workload c1 --skip-cpu > gpurun_out/bench_c1.log 2>&1; tail -n 1 gpurun_out/bench_c1.log | cut -c1-400
timeout 300 python bench.py --skip-cpu --steps 20 > gpurun_out/bench_c2.log 2>&1; tail -n 1 gpurun_out/bench_c2.log | cut -c1-300
