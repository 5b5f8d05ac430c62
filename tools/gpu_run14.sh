#!/bin/bash
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 600 python -m pytest tests/test_gpu_sharded.py -m gpu -x -q 2>&1 | tail -15
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/sharded_commit_bench.py 24 256 1 3 > gpurun_out/sharded_$N.json 2> gpurun_out/sharded_$N.err; echo rc=$?
tail -2 gpurun_out/sharded_$N.json; grep -v "^\*\|OMP_NUM\|^$" gpurun_out/sharded_$N.err | tail -12
