#!/bin/bash
# Single-GPU trip: parity tests, smoke, default bench line (+ reference arm), per-kernel times (CUPTI).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1500 python -m pytest tests -q -m gpu --timeout=300 -x > gpurun_out/tests.log 2>&1; echo "tests exit $?" >> gpurun_out/tests.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
for w in c2 c4 c1; do for d in gaussian trained; do
  timeout 200 python tools/kernel_times.py $w --data $d --pdl 0 > gpurun_out/kt0_${w}_${d}.log 2>&1
done; done
timeout 200 python tools/kernel_times.py c2 --dtype fp16 --pdl 0 > gpurun_out/kt0_c2_fp16.log 2>&1
timeout 200 python tools/kernel_times.py c2 --cf 1024 --pdl 0 > gpurun_out/kt0_c2_cf.log 2>&1
timeout 200 python tools/kernel_times.py c4 --cf 1024 --pdl 0 > gpurun_out/kt0_c4_cf.log 2>&1
timeout 900 python bench.py > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench exit $?" >> gpurun_out/bench.log
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.log 2>&1; echo "ref exit $?" >> gpurun_out/bench_ref.log
for f in tests smoke bench bench_ref; do echo "== $f"; tail -n 3 gpurun_out/$f.log | cut -c1-600; done
tail -n 5 gpurun_out/bench.err
grep -h "FAILED\|Error" gpurun_out/tests.log | head -20
for f in gpurun_out/kt0_*.log; do echo "== $f"; grep -v "Warn\|_warn" $f | head -8 | cut -c1-160; done
