#!/bin/bash
# mixed integer / FP64 leaf hash: tests, full bench line, launch list, ncu --set full of the new kernel
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest10.log 2>&1; echo pytest rc=$?; tail -3 gpurun_out/pytest10.log
timeout 600 python bench.py > gpurun_out/bench10.json 2> gpurun_out/bench10.err; echo bench rc=$?
python - <<PY
import json
d=json.loads(open('gpurun_out/bench10.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['e2e'], d['stage_ms_per_step'], d['segment_proof']['device_ms_per_proof'], d['segment_proof']['segments_per_s'])
PY
CMD="python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu --no-proof"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:leaf_mixed -c 1 -o gpurun_out/prof_leafmix_r1g -f $CMD > gpurun_out/ncu10.log 2>&1; echo ncu rc=$?
ls -la gpurun_out/*.ncu-rep
