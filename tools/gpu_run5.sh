#!/bin/bash
# launch list of the segment-proof leg (small commit workload so the proof kernels dominate the list)
mkdir -p gpurun_out
CMD="python bench.py --log-height 12 --steps 1 --warmup 1 --no-e2e --no-cpu"
$CMD > gpurun_out/plain5.log 2>&1 || { echo plain failed; tail -5 gpurun_out/plain5.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches_proof_r1d.csv $CMD > gpurun_out/ncu5.log 2>&1
echo ncu rc=$?
wc -l gpurun_out/launches_proof_r1d.csv
