#!/bin/bash
# scratch trip: pair (plain remote arrive) vs single on the named workloads, then the parity tests with pair on
mkdir -p gpurun_out; rm -f gpurun_out/kern.log
run() { echo "== $*" | tee -a gpurun_out/kern.log; env "$@" timeout 300 python tools/bench_kernels.py c2 c4 c1 c3 c5 2>&1 | grep -v "^vq_assign" | tail -n 5 | cut -c1-75 | tee -a gpurun_out/kern.log; }
run VQ_ASSIGN_PAIR=1
run VQ_ASSIGN_PAIR=0
run VQ_ASSIGN_PAIR=1 VQ_ASSIGN_DBG=2
timeout 900 python -m pytest tests -q -m gpu -x --timeout=300 > gpurun_out/tests.log 2>&1; echo "tests exit $?" >> gpurun_out/tests.log
tail -n 4 gpurun_out/tests.log
