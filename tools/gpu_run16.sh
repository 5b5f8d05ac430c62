#!/bin/bash
# r1j evidence at HEAD: gpu tests, full bench line + reference arm, log_blowup 2, proof shapes, launch list of one default step
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest16.log 2>&1; echo pytest rc=$?; tail -3 gpurun_out/pytest16.log
timeout 600 python bench.py > gpurun_out/r1j_bench_full.json 2> gpurun_out/bench16.err; echo bench rc=$?
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1j_bench_reference.json 2> gpurun_out/bench16_ref.err; echo ref rc=$?
timeout 600 python bench.py --log-blowup 2 --no-proof --no-cpu --steps 3 --warmup 3 > gpurun_out/r1j_bench_b2.json 2> gpurun_out/bench16_b2.err; echo b2 rc=$?
python - <<PY
import json
for f in ('r1j_bench_full','r1j_bench_b2'):
    d=json.loads(open('gpurun_out/%s.json'%f).read().strip().splitlines()[-1])
    print(f, d['ms_per_step'], d['value'], d['e2e'] and d['e2e'].get('ms_per_step'), d['stage_ms_per_step'], d.get('segment_proof') and (d['segment_proof']['device_ms_per_proof'], d['segment_proof']['segments_per_s']))
PY
B2S_PROVE_TRACE=1 timeout 600 python tools/leaf_shape.py 0 3 > gpurun_out/r1j_leaf_shape.txt 2>&1; grep -E "proof 2|accepted" gpurun_out/r1j_leaf_shape.txt; grep "reduced openings" gpurun_out/r1j_leaf_shape.txt | tail -1
timeout 600 python tools/keccak_shape.py > gpurun_out/r1j_keccak_shape.txt 2>&1; tail -2 gpurun_out/r1j_keccak_shape.txt
CMD="python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file gpurun_out/r1j_launches.csv $CMD > gpurun_out/ncu16.log 2>&1
echo ncu rc=$?; wc -l gpurun_out/r1j_launches.csv
