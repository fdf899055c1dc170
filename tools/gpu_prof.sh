#!/bin/bash
# ncu --set full of the hot kernels on one workload:  bash tools/gpu_prof.sh c2 [kernel-regex]
w=${1:-c2}; rx=${2:-"assign_tc_kernel|pass_kernel"}
mkdir -p gpurun_out
timeout 300 python tools/prof_kernels.py $w 3 > gpurun_out/prof_plain_$w.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"$rx" -s 2 -c 2 -f -o gpurun_out/prof_$w python tools/prof_kernels.py $w 3 > gpurun_out/ncu_$w.log 2>&1
echo "ncu $w exit $?" >> gpurun_out/ncu_$w.log
tail -n 3 gpurun_out/ncu_$w.log
