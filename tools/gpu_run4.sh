#!/bin/bash
# full default bench + reference arm + smoke; outputs under gpurun_out/
set -x
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke_r1d.log 2>&1; echo "smoke rc=$?"
python bench.py > gpurun_out/bench_r1d.json 2> gpurun_out/bench_r1d.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/bench_r1d.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_r1d.json 2> gpurun_out/bench_ref_r1d.err; echo "ref rc=$?"
tail -c 1500 gpurun_out/bench_ref_r1d.json
tail -5 gpurun_out/bench_r1d.err gpurun_out/smoke_r1d.log
