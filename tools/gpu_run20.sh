#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_bn254.py -x -q > gpurun_out/pytest20.log 2>&1; echo pytest rc=$?; tail -40 gpurun_out/pytest20.log
