#!/bin/bash
mkdir -p gpurun_out
run() {
  env $1 timeout 300 python bench.py --no-e2e --no-cpu --no-proof --steps 2 --warmup 1 > gpurun_out/b9.json 2> gpurun_out/b9.err
  echo "$1 rc=$? $(python - <<PY
import json
d=json.loads(open('gpurun_out/b9.json').read().strip().splitlines()[-1])
print(d['stage_ms_per_step']['leaf_hash_ms'], d['root'][:1])
PY
)"
}
for sp in 8,2 7,2 6,3 5,4 4,5 3,5 8,0 0,8; do run "B2S_HASH_VARIANT=5 B2S_HASH_SPLIT=$sp"; done
