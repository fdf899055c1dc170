#!/bin/bash
# scratch trip: new tests, kNN timing, c3 bench line
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x --timeout=300 > gpurun_out/tests.log 2>&1; echo "tests exit $?" >> gpurun_out/tests.log
tail -n 12 gpurun_out/tests.log
timeout 300 python tools/bench_knn.py > gpurun_out/knn_bench.log 2>&1; echo "knn exit $?" >> gpurun_out/knn_bench.log
timeout 300 python tools/bench_knn.py --patches 32 --anchors 256 > gpurun_out/knn_bench_8k.log 2>&1
timeout 300 python tools/bench_knn.py --patches 16 --anchors 256 --channels 128 > gpurun_out/knn_bench_c128.log 2>&1
timeout 300 python bench.py --workload c3 --skip-cpu > gpurun_out/bench_c3.log 2>&1; echo "c3 exit $?" >> gpurun_out/bench_c3.log
for f in knn_bench knn_bench_8k knn_bench_c128 bench_c3; do echo "== $f"; tail -n 3 gpurun_out/$f.log | cut -c1-1800; done
