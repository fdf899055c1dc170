#!/bin/bash
# verification trip: parity tests, smoke, default bench line, kernel timings
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu --timeout=300 > gpurun_out/tests.log 2>&1; echo "tests exit $?" >> gpurun_out/tests.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_c2.log 2>&1; echo "bench exit $?" >> gpurun_out/bench_c2.log
timeout 300 python tools/bench_kernels.py c2 c4 c1 > gpurun_out/kern.log 2>&1
for f in tests smoke bench_c2 kern; do echo "== $f"; tail -n 6 gpurun_out/$f.log | cut -c1-1800; done
