#!/bin/bash
# Final N-GPU check: DP correctness, the default bench line (C4 + C2 records, dp_check), C5 train step under DDP.
N=${1:-4}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 200 $TR --master-port 29533 tools/dp_check.py --quick --json gpurun_out/dp_checkF_$N.json > gpurun_out/dp_checkF_$N.log 2>&1; echo "dp_check exit $?" >> gpurun_out/dp_checkF_$N.log
timeout 300 $TR --master-port 29555 bench.py --gpus $N > gpurun_out/benchF_$N.log 2> gpurun_out/benchF_$N.err; echo "bench exit $?" >> gpurun_out/benchF_$N.log
timeout 150 $TR --master-port 29577 tools/bench_train_step.py --batch 2048 --steps 3 --warmup 2 --quantizer native > gpurun_out/train_step_$N.log 2>&1; echo "train exit $?" >> gpurun_out/train_step_$N.log
for f in dp_checkF_$N benchF_$N train_step_$N; do echo "== $f"; grep -v "^\*\*\*\|OMP_NUM" gpurun_out/$f.log | tail -n 5 | cut -c1-600; done
tail -3 gpurun_out/benchF_$N.err
