#!/bin/bash
# Multi-GPU trip (gpurun --gpus N): DP correctness vs the oracle, per-rank kernel timeline (CUPTI), bench line with dp_check.
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
nvidia-smi topo -m > gpurun_out/topo_$N.txt 2>&1
if [ "$N" = "2" ]; then timeout 900 python -m pytest tests/test_gpu_3_multigpu.py tests/test_gpu_6_dtypes.py -q -m gpu --timeout=300 > gpurun_out/tests_mg.log 2>&1; echo "tests exit $?" >> gpurun_out/tests_mg.log; fi
timeout 600 $TR --master-port 29533 tools/dp_check.py --quick --json gpurun_out/dp_check_$N.json > gpurun_out/dp_check_$N.log 2>&1; echo "dp_check exit $?" >> gpurun_out/dp_check_$N.log
for x in peer nccl; do for w in c2 c4; do
  timeout 300 $TR --master-port 29544 tools/kernel_times.py $w --exchange $x --pdl 0 --json gpurun_out/timeline${N}_${w}_$x.json > gpurun_out/timeline${N}_${w}_$x.log 2>&1
done; done
timeout 300 python tools/kernel_times.py c2 --pdl 0 > gpurun_out/timeline1_c2.log 2>&1
timeout 300 python tools/kernel_times.py c4 --pdl 0 > gpurun_out/timeline1_c4.log 2>&1
timeout 600 $TR --master-port 29555 bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_$N.log 2> gpurun_out/bench_$N.err; echo "bench exit $?" >> gpurun_out/bench_$N.log
timeout 300 python bench.py --steps 30 --warmup 5 --skip-cpu --no-extras > gpurun_out/bench_1.log 2> gpurun_out/bench_1.err; echo "bench exit $?" >> gpurun_out/bench_1.log
[ "$N" = "2" ] && timeout 200 $TR --master-port 29566 tools/symm_probe.py > gpurun_out/symm_probe_$N.log 2>&1; echo "probe exit $?" >> gpurun_out/symm_probe_$N.log
for f in tests_mg dp_check_$N bench_$N bench_1 symm_probe_$N; do echo "== $f"; tail -n 6 gpurun_out/$f.log 2>/dev/null | cut -c1-1200; done
tail -3 gpurun_out/bench_$N.err
for f in gpurun_out/timeline*_*.log; do echo "== $f"; grep -v "Warn\|_warn\|^rank" $f | tail -14 | cut -c1-200; done
