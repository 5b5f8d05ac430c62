#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_bn254.py -x -q > gpurun_out/pytest19.log 2>&1; echo pytest rc=$?; tail -30 gpurun_out/pytest19.log
