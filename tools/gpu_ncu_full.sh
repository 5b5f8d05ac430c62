#!/bin/bash
# full-size (2^24 x 256) ncu captures for profiles/: NTT passes of the first column group, then the leaf-hash launch
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu --no-proof"
$CMD > gpurun_out/plain_full_r1f.log 2>&1 || { echo plain failed; tail -5 gpurun_out/plain_full_r1f.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:ntt_ -c 6 -o gpurun_out/prof_ntt_r1f -f $CMD > gpurun_out/ncu_ntt_r1f.log 2>&1
echo ncu ntt rc=$?
