#!/bin/bash
# r1l: ncu --set full of the dominant kernel at HEAD (one leaf_mixed_kernel launch of the full-size commit), after the same
# command exited 0 without ncu
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu --no-proof"
$CMD > gpurun_out/plain_r1l.log 2>&1 || { echo plain failed; tail -5 gpurun_out/plain_r1l.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:leaf_mixed -c 1 -o gpurun_out/prof_leaf_r1l -f $CMD > gpurun_out/ncu_leaf_r1l.log 2>&1
echo ncu rc=$?; ls -la gpurun_out/prof_leaf_r1l.ncu-rep
