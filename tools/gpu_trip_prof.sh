#!/bin/bash
# Single-GPU trip + profiles: tools/gpu_trip.sh, then the ncu launch list of the bench command and ncu --set full of the
# hot kernels at C2 / C4 (each ncu pass only after the same command exited 0 without ncu).
bash tools/gpu_trip.sh
B="python bench.py --steps 5 --warmup 3 --no-extras --skip-cpu"
timeout 300 $B > gpurun_out/bench_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_bench.csv $B > gpurun_out/ncu_bench.log 2>&1
echo "ncu launch list exit $?"
bash tools/gpu_prof.sh c4
bash tools/gpu_prof.sh c2
ls -la gpurun_out/*.ncu-rep
