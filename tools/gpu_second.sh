#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_1_kernels.py -q -m gpu --timeout=300 > gpurun_out/t1.log 2>&1; echo "t1 exit $?" >> gpurun_out/t1.log
timeout 600 python bench.py --steps 10 --warmup 3 --workload c4 --skip-cpu > gpurun_out/bench_c4.log 2>&1; echo "bench exit $?" >> gpurun_out/bench_c4.log
timeout 300 python bench.py --steps 2 --warmup 3 --skip-cpu > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c2.csv python bench.py --steps 2 --warmup 3 --skip-cpu > gpurun_out/ncu1.log 2>&1
echo "ncu1 exit $?" >> gpurun_out/ncu1.log
timeout 300 python bench.py --steps 2 --warmup 3 --skip-cpu --workload c4 > gpurun_out/plain4.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c4.csv python bench.py --steps 2 --warmup 3 --skip-cpu --workload c4 > gpurun_out/ncu2.log 2>&1
echo "ncu2 exit $?" >> gpurun_out/ncu2.log
for f in t1 bench_c4 ncu1 ncu2; do echo "== $f"; tail -n 4 gpurun_out/$f.log; done
