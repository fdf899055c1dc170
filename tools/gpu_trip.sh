#!/bin/bash
# Round-2 GPU trip: parity tests, smoke, the default bench line (+ reference arm), per-kernel times (CUPTI, PDL off) for
# the arrival-ordered (presort) and block-ordered (sort) visiting orders, fp16 / fp32-tier step times.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1500 python -m pytest tests -q -m gpu --timeout=300 > gpurun_out/tests.log 2>&1; echo "tests exit $?" >> gpurun_out/tests.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench exit $?" >> gpurun_out/bench.log
for w in c2 c4; do for d in gaussian trained; do for m in presort sort; do
  timeout 200 python tools/kernel_times.py $w --data $d --mode $m --pdl 0 > gpurun_out/kt0_${w}_${d}_${m}.log 2>&1
done; done; done
for dt in bf16 fp16; do timeout 200 python tools/kernel_times.py c2 --dtype $dt --pdl 0 > gpurun_out/kt0_c2_module_$dt.log 2>&1; done
for w in c2 c4; do for s in fp32 fp32_simt; do timeout 300 python tools/kernel_times.py $w --dtype fp32 --search $s --pdl 0 --steps 5 > gpurun_out/kt0_${w}_$s.log 2>&1; done; done
for f in tests smoke bench; do echo "== $f"; tail -n 3 gpurun_out/$f.log | cut -c1-2500; done
tail -n 5 gpurun_out/bench.err
grep -h "FAILED\|Error\|error" gpurun_out/tests.log | head -30
for f in gpurun_out/kt0_*.log; do echo "== $f"; grep -v Warn $f | head -9 | cut -c1-160; done
