#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/bench_eager_gpu.py c2 c4 c1 > gpurun_out/eager_gpu.log 2>&1; echo "exit $?" >> gpurun_out/eager_gpu.log; cat gpurun_out/eager_gpu.log
timeout 600 python -m pytest tests/test_gpu_2_tensorcore.py tests/test_gpu_4_configs.py -q -x --timeout=300 2>&1 | tail -n 2
timeout 300 python bench.py --skip-cpu --steps 20 2>&1 | tail -n 1 | cut -c1-250
