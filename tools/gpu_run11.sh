#!/bin/bash
mkdir -p gpurun_out
timeout 600 python bench.py --no-cpu --no-e2e --no-proof --steps 3 > gpurun_out/bench11.json 2> gpurun_out/bench11.err; echo bench rc=$?
python - <<PY
import json
d=json.loads(open('gpurun_out/bench11.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['stage_ms_per_step'])
PY
timeout 600 python tools/leaf_shape.py 0 3 2>&1 | tail -3
timeout 600 python tools/keccak_shape.py 2>&1 | tail -2
