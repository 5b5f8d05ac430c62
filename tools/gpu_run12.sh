#!/bin/bash
mkdir -p gpurun_out
for L in 15 16 17 18 19 20; do
  echo "== MIN_LOG $L"
  B2S_HASH_MIN_LOG=$L timeout 600 python tools/leaf_shape.py 0 3 2>&1 | grep "proof 2"
  B2S_HASH_MIN_LOG=$L timeout 600 python tools/keccak_shape.py 2>&1 | grep "proof 2"
  B2S_HASH_MIN_LOG=$L timeout 600 python bench.py --log-height 12 --steps 1 --warmup 1 --no-cpu --no-e2e > gpurun_out/b12.json 2>/dev/null
  python - <<PY
import json
d=json.loads(open('gpurun_out/b12.json').read().strip().splitlines()[-1])
print('segment', d['segment_proof']['device_ms_per_proof'], d['segment_proof']['segments_per_s'])
PY
done
