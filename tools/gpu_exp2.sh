#!/bin/bash
mkdir -p gpurun_out
for w in c2 c4; do
for p in 74 66 60 52; do
for c in 4 8; do
  VQ_ASSIGN_MAX_PAIRS=$p timeout 120 python tools/overlap_proto.py $w $c 2>&1 | tail -n 1 | tee -a gpurun_out/overlap.log
done; done; done
