#!/bin/bash
mkdir -p gpurun_out
python tools/keccak_shape.py > gpurun_out/keccak_plain.log 2>&1 || { echo plain failed; tail gpurun_out/keccak_plain.log; exit 1; }
tail -3 gpurun_out/keccak_plain.log
ncu --set full --clock-control none --import-source on -k regex:quotient_kernel -c 1 -o gpurun_out/prof_quot_r1e -f python tools/keccak_shape.py > gpurun_out/ncu_quot.log 2>&1
echo rc=$?
