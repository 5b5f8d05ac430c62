#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu --no-proof"
$CMD > gpurun_out/plain_full2_r1e.log 2>&1 || { echo plain failed; tail -5 gpurun_out/plain_full2_r1e.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:tree_level_kernel -c 1 -o gpurun_out/prof_hash_r1e -f $CMD > gpurun_out/ncu_hash_r1e.log 2>&1
echo ncu hash rc=$?
