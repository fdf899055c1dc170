#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_5_presort.py -q -m gpu -x --timeout=300 2>&1 | tail -n 25
