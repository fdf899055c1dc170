#!/bin/bash
# A/B on ONE box: search kernel of other trees (ab_*/, built in place) vs the current tree.
mkdir -p gpurun_out
export VQ_BENCH_SKIP_PASS=1
TREES=${TREES:-"ab_6adf482 . ab_vA ab_vB ab_vC"}
for rep in 1 2; do
  for t in $TREES; do
    n=$(basename $t); [ "$t" = "." ] && n=head
    (cd $t && timeout 100 python tools/bench_kernels.py ${SHAPES:-c4}) > gpurun_out/ab_${n}_$rep.log 2>&1
    echo "$n rep $rep: $(grep -v Warn gpurun_out/ab_${n}_$rep.log | tail -2 | tr '\n' ' ')"
  done
done
