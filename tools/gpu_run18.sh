#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest18.log 2>&1; echo pytest rc=$?; tail -3 gpurun_out/pytest18.log
for s in "" 1; do
  echo "== B2S_HOST_SCALAR=$s"
  if [ -n "$s" ]; then export B2S_HOST_SCALAR=1; else unset B2S_HOST_SCALAR; fi
  B2S_PROVE_TRACE=1 timeout 600 python tools/keccak_shape.py 2>&1 | grep -E "observe opened|ms|accepted" | tail -6
  timeout 600 python tools/keccak_shape.py 2>&1 | tail -2
  timeout 600 python bench.py --log-height 12 --steps 3 --warmup 1 --no-cpu --no-e2e > gpurun_out/b18_$s.json 2>/dev/null
  python - <<PY
import json
d=json.loads(open('gpurun_out/b18_$s.json').read().strip().splitlines()[-1])
print('segment', d['segment_proof']['device_ms_per_proof'], d['segment_proof']['segments_per_s'], d['segment_proof']['stage_ms'])
PY
done
