#!/bin/bash
# leaf hash: number of integer warps per 8-warp block (the rest run the FP64 code path), full-size commit
for k in 2 3 4 5 6; do
  B2S_HASH_F64=$k python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu --no-proof 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('int warps $k:', round(d['ms_per_step'],1), 'ms; leaf', round(d['stage_ms_per_step']['leaf_hash_ms'],1))"
done
