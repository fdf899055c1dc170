#!/bin/bash
mkdir -p gpurun_out
export VQ_B200_LIB=$PWD/vq-vae_b200/libvq_b200_s4r4.so
for w in c2 c4; do
timeout 300 python tools/prof_kernels.py $w 3 > gpurun_out/plain_$w.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/l_$w.csv python tools/prof_kernels.py $w 3 > gpurun_out/ncu_l_$w.log 2>&1
echo "$w exit $?"
done
