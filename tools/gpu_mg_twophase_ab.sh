#!/bin/bash
# Two-phase exchange A/B at N GPUs: DP correctness, per-rank kernel tables with VQ_PEER_TWO_PHASE = 0 / 1, bench.
N=${1:-4}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29533 tools/dp_check.py --quick --json gpurun_out/dp_check2p_$N.json > gpurun_out/dp_check2p_$N.log 2>&1; echo "dp_check exit $?" >> gpurun_out/dp_check2p_$N.log
for w in c4 c2; do
  for tp in 0 1; do
    VQ_PEER_TWO_PHASE=$tp timeout 300 $TR --master-port 2954$tp tools/kernel_times.py $w --exchange peer --pdl 0 > gpurun_out/timeline${N}_${w}_tp$tp.log 2>&1
    VQ_PEER_TWO_PHASE=$tp timeout 300 $TR --master-port 2956$tp tools/kernel_times.py $w --exchange peer > gpurun_out/timeline${N}_${w}_tp${tp}_pdl.log 2>&1
  done
done
timeout 600 $TR --master-port 29555 bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench2p_$N.log 2> gpurun_out/bench2p_$N.err; echo "bench exit $?" >> gpurun_out/bench2p_$N.log
for f in dp_check2p_$N bench2p_$N; do echo "== $f"; grep -v "^\*\*\*\|OMP_NUM" gpurun_out/$f.log | tail -n 6 | cut -c1-700; done
tail -3 gpurun_out/bench2p_$N.err
for f in gpurun_out/timeline${N}_*tp*.log; do echo "== $f"; grep -v "Warn\|_warn\|^rank" $f | tail -12 | cut -c1-260; done
