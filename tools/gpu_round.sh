#!/bin/bash
# One GPU trip: parity tests, smoke, bench (c2 default + c4), ncu launch list of the bench command, ncu --set full
# captures of the two hot kernels.  Everything bounded by its own timeout; logs land in gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1200 python -m pytest tests -q -m gpu --timeout=300 > gpurun_out/tests.log 2>&1; echo "tests exit $?" >> gpurun_out/tests.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_c2.log 2>&1; echo "bench exit $?" >> gpurun_out/bench_c2.log
timeout 600 python bench.py --workload c4 --skip-cpu > gpurun_out/bench_c4.log 2>&1; echo "bench exit $?" >> gpurun_out/bench_c4.log
for w in c1 c3 c5; do timeout 600 python bench.py --workload $w --skip-cpu > gpurun_out/bench_$w.log 2>&1; echo "bench exit $?" >> gpurun_out/bench_$w.log; done
for w in c2 c4; do timeout 600 python bench.py --workload $w --skip-cpu --data trained > gpurun_out/bench_${w}_trained.log 2>&1; echo "bench exit $?" >> gpurun_out/bench_${w}_trained.log; done
timeout 300 python tools/bench_kernels.py c2 c4 c1 c5 > gpurun_out/kern.log 2>&1
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.log 2>&1; echo "ref exit $?" >> gpurun_out/bench_ref.log
# launch list of the default bench command (cold-cache, serialised: shares only)
timeout 300 python bench.py --steps 3 --warmup 3 --skip-cpu > gpurun_out/plain_c2.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_c2.csv python bench.py --steps 3 --warmup 3 --skip-cpu > gpurun_out/ncu_l2.log 2>&1
echo "ncu launches c2 exit $?" >> gpurun_out/ncu_l2.log
timeout 300 python bench.py --steps 3 --warmup 3 --skip-cpu --workload c4 > gpurun_out/plain_c4.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_c4.csv python bench.py --steps 3 --warmup 3 --skip-cpu --workload c4 > gpurun_out/ncu_l4.log 2>&1
echo "ncu launches c4 exit $?" >> gpurun_out/ncu_l4.log
# full captures of the two hot kernels
for w in c4 c2; do
  timeout 300 python tools/prof_kernels.py $w 3 > gpurun_out/prof_plain_$w.log 2>&1 &&
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"assign_tc_kernel|pass_kernel|scatter_ranked" -s 3 -c 6 -f -o gpurun_out/prof_$w python tools/prof_kernels.py $w 3 > gpurun_out/ncu_$w.log 2>&1
  echo "ncu $w exit $?" >> gpurun_out/ncu_$w.log
done
for f in tests smoke bench_c2 bench_c4 bench_ref ncu_l2 ncu_l4 ncu_c4 ncu_c2; do echo "== $f"; tail -n 3 gpurun_out/$f.log | cut -c1-1500; done
