#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x --timeout=300 2>&1 | tail -n 2
timeout 300 python tools/bench_kernels.py c2 c4 c1 2>&1 | grep "stand-alone" | tee -a gpurun_out/exp2.log
timeout 600 python bench.py --skip-cpu > gpurun_out/bench_c2.log 2>&1; tail -n 1 gpurun_out/bench_c2.log | cut -c1-400
timeout 600 python bench.py --skip-cpu --workload c4 > gpurun_out/bench_c4.log 2>&1; tail -n 1 gpurun_out/bench_c4.log | cut -c1-400
timeout 600 python bench.py --skip-cpu --workload c1 > gpurun_out/bench_c1.log 2>&1; tail -n 1 gpurun_out/bench_c1.log | cut -c1-400
