#!/bin/bash
# round-1 validation of HEAD: gpu tests, full bench line, reference arm
mkdir -p gpurun_out
nproc > gpurun_out/host.txt; free -g >> gpurun_out/host.txt; nvidia-smi -L >> gpurun_out/host.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest6.log 2>&1; echo pytest rc=$? | tee -a gpurun_out/pytest6.log
tail -5 gpurun_out/pytest6.log
timeout 600 python bench.py > gpurun_out/bench6.json 2> gpurun_out/bench6.err; echo bench rc=$?
tail -c 1500 gpurun_out/bench6.json
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench6_ref.json 2> gpurun_out/bench6_ref.err; echo ref rc=$?
