#!/bin/bash
# A/B of the mixed integer / FP64 leaf hash: B2S_HASH_F64 = off (old kernels) or the number of integer warps of 8
mkdir -p gpurun_out
timeout 100 tools/bin/p2_fp64 > gpurun_out/p2_fp64c.txt 2>&1; head -3 gpurun_out/p2_fp64c.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest8.log 2>&1; echo pytest rc=$?; tail -3 gpurun_out/pytest8.log
for k in off 8 6 5 4 3; do
  B2S_HASH_F64=$k timeout 300 python bench.py --no-e2e --no-cpu --no-proof --steps 2 --warmup 1 > gpurun_out/bench8_$k.json 2> gpurun_out/bench8_$k.err
  echo "F64=$k rc=$? $(python - <<PY
import json
d=json.loads(open('gpurun_out/bench8_$k.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['stage_ms_per_step'], d['root'][:2])
PY
)"
done
