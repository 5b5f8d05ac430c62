for ge in 1048576 4194304 16777216 67108864; do
B2S_LDE_GROUP_ELEMS=$ge python bench.py --steps 3 --warmup 2 --log-height 20 --no-e2e --no-cpu 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$ge', d['stage_ms_per_step']['lde_ms'])"
done
python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu --log-height 20 > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:ntt_ -s 8 -c 4 -o gpurun_out/prof_ntt_r1b python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu --log-height 20 > gpurun_out/ncu3.log 2>&1
echo rc=$?
