#!/bin/bash
# 8-GPU confirmation with tight time bounds (charged 8x): symmetric-memory probe first; if it fails the rest runs on the
# CUDA-IPC exchange (VQ_PEER_SYMM=0).  Then the default bench, DP correctness, per-rank kernel tables.
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 100 $TR --master-port 29566 tools/symm_probe.py > gpurun_out/symm_probe_$N.log 2>&1; rc=$?; echo "probe exit $rc" >> gpurun_out/symm_probe_$N.log
if [ $rc -ne 0 ]; then export VQ_PEER_SYMM=0; echo "SYMM PROBE FAILED rc=$rc: using VQ_PEER_SYMM=0"; fi
timeout 200 $TR --master-port 29555 bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench2p_$N.log 2> gpurun_out/bench2p_$N.err; echo "bench exit $?" >> gpurun_out/bench2p_$N.log
timeout 150 $TR --master-port 29533 tools/dp_check.py --quick --json gpurun_out/dp_check2p_$N.json > gpurun_out/dp_check2p_$N.log 2>&1; echo "dp_check exit $?" >> gpurun_out/dp_check2p_$N.log
for w in c4 c2; do
  timeout 90 $TR --master-port 29544 tools/kernel_times.py $w --exchange peer --pdl 0 --json gpurun_out/timeline${N}_${w}_mc.json > gpurun_out/timeline${N}_${w}_mc.log 2>&1
done
nvidia-smi topo -m > gpurun_out/topo_$N.txt 2>&1
for f in symm_probe_$N dp_check2p_$N bench2p_$N; do echo "== $f"; grep -v "^\*\*\*\|OMP_NUM" gpurun_out/$f.log | tail -n 6 | cut -c1-700; done
tail -3 gpurun_out/bench2p_$N.err
for f in gpurun_out/timeline${N}_*_mc.log; do echo "== $f"; grep -v "Warn\|_warn\|^rank" $f | tail -12 | cut -c1-260; done
