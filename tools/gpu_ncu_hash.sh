#!/bin/bash
# one full ncu capture of the leaf-hash kernel (2^20 x 256 commit); $1 = tag
TAG=${1:-x}
mkdir -p gpurun_out
python tools/p2_sweep.py > gpurun_out/p2_plain_$TAG.log 2>&1 || { echo plain failed; exit 1; }
cat gpurun_out/p2_plain_$TAG.log
ncu --set full --clock-control none --import-source on -k regex:tree_level_kernel -s 2 -c 1 -o gpurun_out/prof_hash_$TAG -f python tools/p2_sweep.py > gpurun_out/ncu_hash_$TAG.log 2>&1
echo ncu rc=$?
