#!/bin/bash
mkdir -p gpurun_out
for w in c3 c5 c1; do
timeout 600 python bench.py --workload $w --skip-cpu $( [ $w = c5 ] && echo "--steps 10" ) > gpurun_out/bench_$w.log 2>&1; echo "exit $?" >> gpurun_out/bench_$w.log
echo "== $w"; tail -n 2 gpurun_out/bench_$w.log | cut -c1-1500
done
