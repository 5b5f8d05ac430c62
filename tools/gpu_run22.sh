#!/bin/bash
# final evidence of the round at HEAD, as the driver runs it: gpu tests, smoke, default bench line, reference arm
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest22.log 2>&1; echo pytest rc=$?; tail -3 gpurun_out/pytest22.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py > gpurun_out/r1n_bench_full.json 2> gpurun_out/bench22.err; echo bench rc=$?
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1n_bench_reference.json 2> gpurun_out/bench22_ref.err; echo ref rc=$?
python - <<PY
import json
d=json.loads(open('gpurun_out/r1n_bench_full.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['e2e'] and d['e2e'].get('ms_per_step'), d.get('segment_proof') and (d['segment_proof']['device_ms_per_proof'], d['segment_proof']['segments_per_s']), d['clocks'])
r=json.loads(open('gpurun_out/r1n_bench_reference.json').read().strip().splitlines()[-1])
print('reference arm', r['value'], r['unit'], r['cpu_baseline']['cores'])
PY
