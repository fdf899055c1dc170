#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x --timeout=300 > gpurun_out/tests.log 2>&1; echo "tests exit $?" >> gpurun_out/tests.log
tail -n 12 gpurun_out/tests.log
timeout 600 python bench.py --skip-cpu > gpurun_out/bench_c2.log 2>&1; echo "bench exit $?" >> gpurun_out/bench_c2.log
timeout 600 python bench.py --workload c4 --skip-cpu > gpurun_out/bench_c4.log 2>&1; echo "bench exit $?" >> gpurun_out/bench_c4.log
for f in bench_c2 bench_c4; do echo "== $f"; tail -n 3 gpurun_out/$f.log | cut -c1-1900; done
