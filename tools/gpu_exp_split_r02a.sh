#!/bin/bash
# Experiment trip (1 GPU): search epilogue split (tuning build, wrong-result timing switches), TMEM read bandwidth,
# gather pass with window-local sorting, C5 train step.
mkdir -p gpurun_out
timeout 120 tools/micro/ldtm_bw > gpurun_out/ldtm_bw.log 2>&1
for dbg in 0 1 2 4; do
  echo "== VQ_ASSIGN_DBG=$dbg (0 = real kernel; 1 = TMEM read, no fold; 2 = no TMEM read; 4 = leader-only epilogue)" >> gpurun_out/search_split.log
  VQ_B200_LIB=$PWD/vqvae_b200/libvq_b200_tune.so VQ_ASSIGN_DBG=$dbg VQ_BENCH_SKIP_PASS=1 timeout 200 python tools/bench_kernels.py c2 c4 >> gpurun_out/search_split.log 2>&1
done
for m in 0 1 2 3; do
  echo "== VQ_ASSIGN_MODE=$m" >> gpurun_out/search_split.log
  VQ_B200_LIB=$PWD/vqvae_b200/libvq_b200_tune.so VQ_ASSIGN_MODE=$m VQ_BENCH_SKIP_PASS=1 timeout 200 python tools/bench_kernels.py c2 >> gpurun_out/search_split.log 2>&1
done
timeout 300 python tools/micro/pass_window.py c2 c4 > gpurun_out/pass_window.log 2>&1
timeout 300 python tools/micro/pass_window.py c2 c4 --data trained >> gpurun_out/pass_window.log 2>&1
timeout 600 python tools/bench_train_step.py --batch 2048 --steps 3 --warmup 2 > gpurun_out/train_step_1.log 2>&1
cat gpurun_out/ldtm_bw.log gpurun_out/search_split.log gpurun_out/pass_window.log | grep -v Warn | cut -c1-220
tail -2 gpurun_out/train_step_1.log | cut -c1-1500
