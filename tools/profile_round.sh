#!/bin/bash
# Profiling recipe of one round (run on the GPU box: gpurun -- 'bash tools/profile_round.sh TAG').
#   1. bench.py without a profiler (the numbers)            -> gpurun_out/bench_TAG.json
#   2. launch list of the same command under ncu            -> gpurun_out/launches_TAG.csv
#   3. one --set full capture per hot kernel, at the bench's shapes (C=65536, D=20, S=64)
# Every step runs under its own timeout; a step only runs after the previous one exited 0.
set -u
TAG=${1:-rXX}
O=gpurun_out
mkdir -p $O
NCU_FULL="ncu --set full --clock-control none --import-source on"
step() { echo "== $*" >&2; timeout 600 "$@"; }

step python bench.py > $O/bench_$TAG.json 2> $O/bench_$TAG.err || { echo "bench failed" >&2; exit 1; }
tail -1 $O/bench_$TAG.json
step ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_$TAG.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_launches_$TAG.log 2>&1 || echo "launch list failed" >&2

for L in 16 128; do
  step $NCU_FULL -k regex:hmc_kernel -s 1 -c 1 -f -o $O/hmc_L${L}_$TAG \
      python tools/perf_hmc.py --samples 64 --Ls $L --eps 0.02 > $O/ncu_hmc_L${L}_$TAG.log 2>&1 || echo "hmc L=$L capture failed" >&2
done
step $NCU_FULL -k regex:nuts_kernel -s 1 -c 1 -f -o $O/nuts_rb_$TAG \
    python tools/perf_hmc.py --sampler nuts --samples 96 --Ls 1 --eps 0.02 > $O/ncu_nuts_$TAG.log 2>&1 || echo "nuts capture failed" >&2
step $NCU_FULL -k regex:rwmh -s 1 -c 1 -f -o $O/rwmh_$TAG \
    python tools/perf_hmc.py --sampler rwmh --target standard_normal --chains 1048576 --samples 64 --Ls 1 --eps 0.5 \
    > $O/ncu_rwmh_$TAG.log 2>&1 || echo "rwmh capture failed" >&2
ls -la $O | tail -20
