#!/bin/bash
# Short multi-GPU check: symmetric-memory probe, DP correctness, per-rank kernel tables with the fused exchange.
N=${1:-4}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29566 tools/symm_probe.py > gpurun_out/symm_probe_$N.log 2>&1; echo "probe exit $?" >> gpurun_out/symm_probe_$N.log
timeout 600 $TR --master-port 29533 tools/dp_check.py --quick --json gpurun_out/dp_check_$N.json > gpurun_out/dp_check_$N.log 2>&1; echo "dp_check exit $?" >> gpurun_out/dp_check_$N.log
for w in c2 c4; do
  timeout 300 $TR --master-port 29544 tools/kernel_times.py $w --exchange peer --pdl 0 --json gpurun_out/timeline${N}_${w}_peer.json > gpurun_out/timeline${N}_${w}_peer.log 2>&1
done
timeout 600 $TR --master-port 29555 bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_$N.log 2> gpurun_out/bench_$N.err; echo "bench exit $?" >> gpurun_out/bench_$N.log
for f in symm_probe_$N dp_check_$N bench_$N; do echo "== $f"; grep -v "^\*\*\*\|OMP_NUM" gpurun_out/$f.log | tail -n 6 | cut -c1-900; done
tail -3 gpurun_out/bench_$N.err
for f in gpurun_out/timeline${N}_*.log; do echo "== $f"; grep -v "Warn\|_warn\|^rank" $f | tail -14 | cut -c1-300; done
