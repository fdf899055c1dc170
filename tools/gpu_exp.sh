#!/bin/bash
mkdir -p gpurun_out
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 tools/dp_check.py --quick 2>&1 | grep -v "^\*\|^W\|OMP_NUM\|^$" | tail -n 8
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --steps 30 --warmup 5 > gpurun_out/bench8_c2.log 2>&1; echo "exit $?" >> gpurun_out/bench8_c2.log
tail -n 2 gpurun_out/bench8_c2.log | cut -c1-330
