#!/bin/bash
# One parameterised GPU-box script (run through gpurun; everything it writes lands in gpurun_out/).
#   tools/gpu.sh tests [TAG] [PYTEST_ARGS...]      GPU test suite
#   tools/gpu.sh smoke                             __graft_entry__.smoke()
#   tools/gpu.sh bench TAG [BENCH_ARGS...]         bench.py line -> gpurun_out/TAG_bench.json
#   tools/gpu.sh reference TAG [BENCH_ARGS...]     bench.py --impl reference
#   tools/gpu.sh launches TAG [BENCH_ARGS...]      ncu launch list (gpu__time_duration) of a bench run, after a plain run exits 0
#   tools/gpu.sh ncu TAG KERNEL_REGEX [COUNT] -- CMD...   ncu --set full of kernels matching the regex, after CMD exits 0 plain
#   tools/gpu.sh env NAME=VALUE... -- CMD...       sweep helper: runs CMD under each assignment in turn
# The r1 summaries under profiles/ name tools/gpu_run*.sh / gpu_ncu*.sh: those one-off wrappers were folded into this file.
set -u
mkdir -p gpurun_out
mode=${1:-tests}; shift || true
case "$mode" in
  tests)
    tag=${1:-tests}; shift || true
    timeout 1200 python -m pytest tests -m gpu -x -q "$@" > gpurun_out/${tag}_pytest.log 2>&1; rc=$?
    echo "pytest rc=$rc"; tail -5 gpurun_out/${tag}_pytest.log; exit $rc ;;
  smoke)
    timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3 ;;
  bench)
    tag=$1; shift
    timeout 1500 python bench.py "$@" > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; rc=$?
    echo "bench rc=$rc"; tail -3 gpurun_out/${tag}_bench.err; tail -c 6000 gpurun_out/${tag}_bench.json; exit $rc ;;
  reference)
    tag=$1; shift
    timeout 900 python bench.py --impl reference "$@" > gpurun_out/${tag}_reference.json 2> gpurun_out/${tag}_reference.err; rc=$?
    echo "reference rc=$rc"; tail -c 2000 gpurun_out/${tag}_reference.json; exit $rc ;;
  launches)
    tag=$1; shift
    python bench.py "$@" > gpurun_out/${tag}_plain.log 2>&1 || { echo plain failed; tail -5 gpurun_out/${tag}_plain.log; exit 1; }
    timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/${tag}_launches.csv \
      python bench.py "$@" > gpurun_out/${tag}_ncu.log 2>&1
    echo "ncu rc=$?"; wc -l gpurun_out/${tag}_launches.csv ;;
  ncu)
    tag=$1; regex=$2; shift 2
    count=1
    if [ "${1:-}" != "--" ]; then count=$1; shift; fi
    shift
    "$@" > gpurun_out/${tag}_plain.log 2>&1 || { echo plain failed; tail -5 gpurun_out/${tag}_plain.log; exit 1; }
    timeout 1500 ncu --set full --clock-control none --import-source on -k "regex:$regex" -c $count -o gpurun_out/${tag} -f \
      "$@" > gpurun_out/${tag}_ncu.log 2>&1
    echo "ncu rc=$?"; ls -la gpurun_out/${tag}.ncu-rep ;;
  env)
    assigns=()
    while [ "${1:-}" != "--" ]; do assigns+=("$1"); shift; done
    shift
    for a in "${assigns[@]}"; do echo "== $a"; env "$a" "$@" 2>&1 | tail -2; done ;;
  *) echo "unknown mode $mode"; exit 2 ;;
esac
