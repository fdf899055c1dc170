#!/bin/bash
# ncu --set full of the pass group on c2 and c4
mkdir -p gpurun_out
for w in c4 c2; do
timeout 300 python tools/prof_kernels.py $w 3 > gpurun_out/prof_plain_$w.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"pass_kernel|scatter_ranked" -s 2 -c 2 -f -o gpurun_out/prof_pass_$w python tools/prof_kernels.py $w 3 > gpurun_out/ncu_pass_$w.log 2>&1
echo "ncu $w exit $?" >> gpurun_out/ncu_pass_$w.log
tail -n 2 gpurun_out/ncu_pass_$w.log
done
