#!/bin/bash
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --sharded --steps 3 --warmup 3 --no-cpu --no-e2e --no-proof > gpurun_out/bench15_sharded_$N.json 2> gpurun_out/bench15_$N.err; echo rc=$?
python - <<PY
import json
d=json.loads(open('gpurun_out/bench15_sharded_$N.json').read().strip().splitlines()[-1])
print(d['n_gpus'], d['ms_per_step'], d['value'], d['sharded_commit'])
PY
grep -v "^\*\|OMP_NUM\|^$" gpurun_out/bench15_$N.err | tail -5
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/sharded_commit_bench.py 24 256 1 3 > gpurun_out/r1i_sharded_$N.json 2> gpurun_out/sharded_$N.err; echo rc=$?
tail -1 gpurun_out/r1i_sharded_$N.json | cut -c1-600
