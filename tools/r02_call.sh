#!/bin/bash
# One GPU-box session of round 2: tools/r02_call.sh TAG [steps...]; every step under its own timeout, outputs in gpurun_out/.
set -u
TAG=${1:-r02a}; shift
STEPS=${*:-"tests bench perf configs ncu"}
O=gpurun_out; mkdir -p $O
NCU_FULL="ncu --set full --clock-control none --import-source on"
for s in $STEPS; do
case $s in
tests)
  timeout 1500 python -m pytest tests -m gpu -q --maxfail=60 -p no:cacheprovider > $O/pytest_$TAG.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest_$TAG.log
  tail -5 $O/pytest_$TAG.log ;;
zz)
  timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > $O/pytest_zz_$TAG.log 2>&1; echo "zz alone rc=$?"; tail -3 $O/pytest_zz_$TAG.log ;;
smoke)
  timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke_$TAG.log 2>&1; echo "smoke rc=$?"; tail -2 $O/smoke_$TAG.log ;;
bench)
  timeout 900 python bench.py > $O/bench_$TAG.json 2> $O/bench_$TAG.err; echo "bench rc=$?"; tail -c 3000 $O/bench_$TAG.json; tail -3 $O/bench_$TAG.err ;;
benchref)
  timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref_$TAG.json 2> $O/bench_ref_$TAG.err; echo "benchref rc=$?"; tail -c 600 $O/bench_ref_$TAG.json ;;
perf)
  { timeout 300 python tools/perf_hmc.py --samples 64 --Ls 16,32,64,128 --eps 0.02
    timeout 300 python tools/perf_hmc.py --sampler grahmc --target neals_funnel --samples 64 --Ls 16,64 --eps 0.05
    for C in 1048576 4194304; do timeout 300 python tools/perf_hmc.py --sampler rwmh --target standard_normal --chains $C --samples 64 --Ls 1 --eps 0.532; done
    for t in standard_normal rosenbrock neals_funnel; do e=0.5; [ $t = rosenbrock ] && e=0.02; [ $t = neals_funnel ] && e=0.1
      timeout 300 python tools/perf_hmc.py --sampler nuts --target $t --samples 96 --Ls 1 --eps $e; done
  } > $O/perf_$TAG.jsonl 2>&1; cat $O/perf_$TAG.jsonl ;;
diag)
  timeout 600 python tools/diag_probe.py > $O/diag_probe_$TAG.txt 2>&1; echo "diag rc=$?"; cat $O/diag_probe_$TAG.txt ;;
configs)
  timeout 1200 python tools/run_configs.py --configs 1,2,3,4,5 --out $O/configs_${TAG}_1gpu.jsonl > $O/configs_$TAG.log 2>&1; echo "configs rc=$?"; tail -25 $O/configs_$TAG.log ;;
ncu)
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_${TAG}_bench.csv \
      python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_launches_$TAG.log 2>&1 || echo "launch list failed"
  timeout 600 $NCU_FULL -k regex:rwmh -s 1 -c 1 -f -o $O/rwmh_$TAG python tools/perf_hmc.py --sampler rwmh --target standard_normal \
      --chains 1048576 --samples 64 --Ls 1 --eps 0.532 > $O/ncu_rwmh_$TAG.log 2>&1 || echo "rwmh capture failed"
  for L in 16 128; do
    timeout 600 $NCU_FULL -k regex:hmc_kernel -s 1 -c 1 -f -o $O/hmc_L${L}_$TAG python tools/perf_hmc.py --samples 64 --Ls $L --eps 0.02 \
        > $O/ncu_hmc_L${L}_$TAG.log 2>&1 || echo "hmc L=$L capture failed"
  done
  timeout 600 $NCU_FULL -k regex:nuts_kernel -s 1 -c 1 -f -o $O/nuts_rb_$TAG python tools/perf_hmc.py --sampler nuts --samples 96 --Ls 1 --eps 0.02 \
      > $O/ncu_nuts_$TAG.log 2>&1 || echo "nuts capture failed" ;;
*) echo "unknown step $s" ;;
esac
done
ls -la $O | tail -30
