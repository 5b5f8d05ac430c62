python tools/lde_sweep.py 22 > gpurun_out/plain3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:ntt_ -s 12 -c 4 -o gpurun_out/prof_ntt_r1c python tools/lde_sweep.py 22 > gpurun_out/ncu3c.log 2>&1
echo rc=$?
python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu --log-height 20 > gpurun_out/plain4.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:tree_level -c 1 -o gpurun_out/prof_hash_r1c python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu --log-height 20 > gpurun_out/ncu4c.log 2>&1
echo rc=$?
