#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/kern.log
run() { echo "== $*" | tee -a gpurun_out/kern.log; env "$@" timeout 300 python tools/bench_kernels.py c2 c4 c3 2>&1 | grep -v "^vq_assign" | tail -n 3 | cut -c60-170 | tee -a gpurun_out/kern.log; }
for v in wide4 wide3 wide3r8; do run VQ_B200_LIB=vq-vae_b200/libvq_b200_$v.so; done
run VQ_B200_LIB=vq-vae_b200/libvq_b200.so
VQ_B200_LIB=vq-vae_b200/libvq_b200_wide3.so timeout 600 python -m pytest tests/test_gpu_1_kernels.py -q -x --timeout=120 2>&1 | tail -n 2
