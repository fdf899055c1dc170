#!/bin/bash
# Search kernel, final form, tuning build: where the time goes (timing switches: results are WRONG) and epilogue warps.
mkdir -p gpurun_out; : > gpurun_out/search_split2.log
export VQ_B200_LIB=$PWD/vqvae_b200/libvq_b200_tune.so VQ_BENCH_SKIP_PASS=1
for dbg in 0 1 2; do
  echo "== VQ_ASSIGN_DBG=$dbg (0 = real kernel; 1 = TMEM read, no fold; 2 = no TMEM read)" >> gpurun_out/search_split2.log
  VQ_ASSIGN_DBG=$dbg timeout 100 python tools/bench_kernels.py c2 c4 2>&1 | grep -v Warn | tail -2 >> gpurun_out/search_split2.log
done
for ew in 4 8 16; do
  echo "== VQ_ASSIGN_EPI_WARPS=$ew" >> gpurun_out/search_split2.log
  VQ_ASSIGN_EPI_WARPS=$ew timeout 100 python tools/bench_kernels.py c2 c4 2>&1 | grep -v Warn | tail -2 >> gpurun_out/search_split2.log
done
cat gpurun_out/search_split2.log
