#!/bin/bash
# r1l evidence at HEAD: gpu tests, memcheck of the new kernels, full bench line + reference arm, root-configuration shape
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest21.log 2>&1; echo pytest rc=$?; tail -3 gpurun_out/pytest21.log
timeout 600 python bench.py > gpurun_out/r1l_bench_full.json 2> gpurun_out/bench21.err; echo bench rc=$?
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1l_bench_reference.json 2> gpurun_out/bench21_ref.err; echo ref rc=$?
python - <<PY
import json
d=json.loads(open('gpurun_out/r1l_bench_full.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['e2e'] and d['e2e'].get('ms_per_step'), d['stage_ms_per_step'], d.get('segment_proof') and (d['segment_proof']['device_ms_per_proof'], d['segment_proof']['segments_per_s']), d['roofline'])
PY
timeout 600 python tools/root_shape.py > gpurun_out/r1l_root_shape.txt 2>&1; tail -4 gpurun_out/r1l_root_shape.txt
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_gpu_bn254.py -x -q -k "device_permutation or mmcs_root or fibonacci_proof_bytes" > gpurun_out/r1l_memcheck.txt 2>&1; echo memcheck rc=$?; tail -4 gpurun_out/r1l_memcheck.txt
