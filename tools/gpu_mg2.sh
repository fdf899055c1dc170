#!/bin/bash
# Short N-GPU check of the exchange: DP correctness, per-rank kernel table at C4, bench.
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29533 tools/dp_check.py --quick --json gpurun_out/dp_check3_$N.json > gpurun_out/dp_check3_$N.log 2>&1; echo "dp_check exit $?" >> gpurun_out/dp_check3_$N.log
timeout 120 $TR --master-port 29544 tools/kernel_times.py c4 --exchange peer --pdl 0 > gpurun_out/timeline${N}_c4_merged.log 2>&1
timeout 300 $TR --master-port 29555 bench.py --gpus $N --steps 30 --warmup 5 --no-extras --skip-cpu > gpurun_out/bench3_$N.log 2> gpurun_out/bench3_$N.err; echo "bench exit $?" >> gpurun_out/bench3_$N.log
for f in dp_check3_$N bench3_$N; do echo "== $f"; grep -v "^\*\*\*\|OMP_NUM" gpurun_out/$f.log | tail -n 6 | cut -c1-500; done
tail -3 gpurun_out/bench3_$N.err
grep "n= \|sum of\|exchange" gpurun_out/timeline${N}_c4_merged.log | cut -c1-200
