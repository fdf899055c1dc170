#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x --timeout=300 2>&1 | tail -n 2
timeout 300 python tools/bench_kernels.py c2 c4 c1 2>&1 | tee -a gpurun_out/exp.log
