#!/bin/bash
# 8-GPU trip: bench at N=8 (c2, c4), kept short
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/bench8_c2.log 2>&1; echo "exit $?" >> gpurun_out/bench8_c2.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus 8 --steps 20 --warmup 5 --workload c4 > gpurun_out/bench8_c4.log 2>&1; echo "exit $?" >> gpurun_out/bench8_c4.log
for f in bench8_c2 bench8_c4; do echo "== $f"; tail -n 2 gpurun_out/$f.log | cut -c1-900; done
