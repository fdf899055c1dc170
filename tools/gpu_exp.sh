#!/bin/bash
# experiment trip: parity tests, then A/B of a library variant (VQ_B200_LIB) on gaussian and trained-like latents
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x --timeout=300 2>&1 | tail -n 2
for rep in 1 2; do for v in "" _noblk; do for w in c2 c4; do for d in gaussian trained; do
  VQ_B200_LIB=$PWD/vqvae_b200/libvq_b200$v.so timeout 300 python bench.py --skip-cpu --workload $w --data $d > gpurun_out/b.log 2>&1
  echo "lib[$v] $w $d $(tail -n 1 gpurun_out/b.log | grep -o '"ms_per_step": [0-9.]*\|"ms_per_launch": [0-9.]*' | tr '\n' ' ')" | tee -a gpurun_out/ab.log
done; done; done; done
