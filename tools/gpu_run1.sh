set -x
python bench.py > gpurun_out/bench_full_r1a.log 2> gpurun_out/bench_full_r1a.err; echo rc=$?
tail -c 3000 gpurun_out/bench_full_r1a.log
python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu --log-height 20 > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_h20_r1a.csv python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu --log-height 20 > gpurun_out/ncu1.log 2>&1
echo rc=$?
ncu --set full --clock-control none --import-source on -k regex:row_hash -c 1 -o gpurun_out/prof_rowhash_r1a python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu --log-height 20 > gpurun_out/ncu2.log 2>&1
echo rc=$?
ncu --set full --clock-control none --import-source on -k regex:ntt_pass -s 12 -c 6 -o gpurun_out/prof_ntt_r1a python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu --log-height 20 > gpurun_out/ncu3.log 2>&1
echo rc=$?
