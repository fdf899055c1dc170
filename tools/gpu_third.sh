#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_1_kernels.py -q -m gpu --timeout=300 > gpurun_out/t1.log 2>&1; echo "t1 exit $?" >> gpurun_out/t1.log
timeout 600 python -m pytest tests/test_gpu_2_tensorcore.py -q -m gpu --timeout=300 > gpurun_out/t2.log 2>&1; echo "t2 exit $?" >> gpurun_out/t2.log
timeout 600 python bench.py --steps 20 --warmup 5 --skip-cpu > gpurun_out/bench_c2.log 2>&1; echo "bench exit $?" >> gpurun_out/bench_c2.log
timeout 600 python bench.py --steps 20 --warmup 5 --workload c4 --skip-cpu > gpurun_out/bench_c4.log 2>&1; echo "bench exit $?" >> gpurun_out/bench_c4.log
timeout 300 python tools/prof_kernels.py c4 3 > gpurun_out/prof_plain_c4.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"assign_tc_kernel|pass_kernel" -s 2 -c 4 -f -o gpurun_out/prof_c4 python tools/prof_kernels.py c4 3 > gpurun_out/ncu_c4.log 2>&1
echo "ncu c4 exit $?" >> gpurun_out/ncu_c4.log
timeout 300 python tools/prof_kernels.py c2 3 > gpurun_out/prof_plain_c2.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"assign_tc_kernel|pass_kernel" -s 2 -c 4 -f -o gpurun_out/prof_c2 python tools/prof_kernels.py c2 3 > gpurun_out/ncu_c2.log 2>&1
echo "ncu c2 exit $?" >> gpurun_out/ncu_c2.log
for f in t1 t2 bench_c2 bench_c4 ncu_c4 ncu_c2; do echo "== $f"; tail -n 3 gpurun_out/$f.log | cut -c1-600; done
