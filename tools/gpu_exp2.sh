#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x --timeout=300 2>&1 | tail -n 3
for v in "" _nobulk; do
  echo "variant [$v]" | tee -a gpurun_out/exp2.log
  VQ_B200_LIB=$PWD/vq-vae_b200/libvq_b200$v.so timeout 300 python tools/bench_kernels.py c2 c4 c1 c5 2>&1 | grep "stand-alone" | tee -a gpurun_out/exp2.log
done
