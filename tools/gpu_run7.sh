#!/bin/bash
# r1g: the remaining BASELINE configs on HEAD -- commit at log_blowup 2, leaf-aggregation shape (config 5), Keccak shape (config 4),
# and the ncu launch list of one default bench step
mkdir -p gpurun_out
timeout 600 python bench.py --log-blowup 2 --no-proof --no-cpu --steps 3 --warmup 3 > gpurun_out/bench7_b2.json 2> gpurun_out/bench7_b2.err; echo b2 rc=$?
tail -c 600 gpurun_out/bench7_b2.json
timeout 900 python tools/leaf_shape.py 0 3 > gpurun_out/leaf_shape7.txt 2>&1; echo leaf rc=$?; tail -8 gpurun_out/leaf_shape7.txt
timeout 600 python tools/keccak_shape.py > gpurun_out/keccak_shape7.txt 2>&1; echo keccak rc=$?; tail -6 gpurun_out/keccak_shape7.txt
CMD="python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file gpurun_out/launches_r1g.csv $CMD > gpurun_out/ncu7.log 2>&1
echo ncu rc=$?; wc -l gpurun_out/launches_r1g.csv
