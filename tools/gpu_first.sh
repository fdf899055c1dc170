#!/bin/bash
# First GPU trip: part-1 tests, part-2 (tcgen05) tests in a separate process, smoke, short bench.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_1_kernels.py -q -m gpu -x --timeout=300 > gpurun_out/t1.log 2>&1; echo "t1 exit $?" >> gpurun_out/t1.log
timeout 900 python -m pytest tests/test_gpu_2_tensorcore.py -q -m gpu --timeout=300 > gpurun_out/t2.log 2>&1; echo "t2 exit $?" >> gpurun_out/t2.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_c2.log 2>&1; echo "bench exit $?" >> gpurun_out/bench_c2.log
tail -5 gpurun_out/t1.log gpurun_out/t2.log gpurun_out/smoke.log gpurun_out/bench_c2.log
