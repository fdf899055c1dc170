#!/bin/bash
# A/B on ONE box: search kernel of other trees (ab_*/, built in place) vs the current tree.
# Make a tree from a commit (ab_*/ is git-ignored but travels with gpurun, built .so included):
#     mkdir ab_<sha> && git archive <sha> | tar -x -C ab_<sha> && (cd ab_<sha> && python -m vqvae_b200.build)
# or copy the working tree, change one thing in csrc/, build.  Then:  TREES="ab_<sha> ." SHAPES="c4 c2" bash tools/gpu_ab_trees.sh
# (this is how the run-time plane count regression of round 2 was found: profiles/r02/search_planes_ab.log).
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
