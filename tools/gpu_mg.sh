#!/bin/bash
# multi-GPU trip: peer-exchange test + bench at the box's GPU count
mkdir -p gpurun_out
G=$(nvidia-smi -L | wc -l)
timeout 600 python -m pytest tests/test_gpu_3_multigpu.py -q -m gpu --timeout=300 2>&1 | tail -n 3
for w in c2 c4; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus $G --workload $w > gpurun_out/bench${G}_$w.log 2>&1
echo "bench$G $w exit $?" >> gpurun_out/bench${G}_$w.log
tail -n 2 gpurun_out/bench${G}_$w.log | cut -c1-700
done
