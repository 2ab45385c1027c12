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
nutsq)
  timeout 150 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "nuts" --timeout 40 -p no:cacheprovider > $O/pytest_nutsq_$TAG.log 2>&1; echo "nuts parity rc=$?"; tail -5 $O/pytest_nutsq_$TAG.log
  timeout 300 python -m pytest --timeout 60 tests/test_fullsize_slices_gpu.py tests/test_edge_cases_gpu.py tests/test_fullsize_gpu.py tests/test_abi_gpu.py tests/test_reference_golden_gpu.py tests/test_dense_gaussian_gpu.py -m gpu -q -k "nuts or NUTS or 3-0 or 3-20 or 3-10 or no_spin or golden or step_api" -p no:cacheprovider > $O/pytest_nutsq2_$TAG.log 2>&1; echo "nuts others rc=$?"; tail -8 $O/pytest_nutsq2_$TAG.log
  { for v in quad thread; do for t in standard_normal rosenbrock neals_funnel; do e=0.5; [ $t = rosenbrock ] && e=0.02; [ $t = neals_funnel ] && e=0.1
      echo "variant=$v"; if [ $v = thread ]; then export MCMC_NUTS_THREAD=1; else unset MCMC_NUTS_THREAD; fi
      timeout 60 python tools/perf_hmc.py --sampler nuts --target $t --samples 96 --Ls 1 --eps $e; done; done; unset MCMC_NUTS_THREAD
  } > $O/nuts_quad_ab_$TAG.txt 2>&1; cat $O/nuts_quad_ab_$TAG.txt ;;
nutsp)
  timeout 200 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "nuts" --timeout 60 -p no:cacheprovider > $O/pytest_nutsp_$TAG.log 2>&1; echo "nuts parity rc=$?"; tail -3 $O/pytest_nutsp_$TAG.log
  { for t in standard_normal rosenbrock neals_funnel; do e=0.5; [ $t = rosenbrock ] && e=0.02; [ $t = neals_funnel ] && e=0.1
      timeout 100 python tools/perf_hmc.py --sampler nuts --target $t --samples 96 --Ls 1 --eps $e; done
    timeout 150 python tools/nuts_latency_probe.py
  } > $O/nuts_perf_$TAG.txt 2>&1; cat $O/nuts_perf_$TAG.txt ;;
abmisc)
  { echo "== dense mat-vec: scalar vs DMMA"; timeout 60 ./tools/micro/dense_dmma_ab.bin
    for v in "" _d5; do echo "== hmc lib variant '$v'"; lib=$PWD/mcmc_b200/csrc/libmcmc_b200$v.so
      MCMC_B200_LIB=$lib timeout 120 python tools/perf_hmc.py --samples 64 --Ls 48,64,96,128 --eps 0.02; done
    for v in "" _rw5 _rw3; do echo "== rwmh lib variant '$v'"; lib=$PWD/mcmc_b200/csrc/libmcmc_b200$v.so; [ -f $lib ] || continue
      MCMC_B200_LIB=$lib timeout 120 python tools/perf_hmc.py --sampler rwmh --target standard_normal --chains 4194304 --samples 64 --Ls 1 --eps 0.532; done
  } > $O/abmisc_$TAG.txt 2>&1; cat $O/abmisc_$TAG.txt ;;
multi)
  N=$(nvidia-smi -L | wc -l); echo "GPUs: $N"
  timeout 600 python -m pytest tests/test_multigpu_gpu.py tests/test_abi_gpu.py -m gpu -q --timeout 300 -p no:cacheprovider > $O/pytest_multi_$TAG.log 2>&1; echo "multi tests rc=$?"; tail -4 $O/pytest_multi_$TAG.log
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > $O/bench_${TAG}_${N}gpu.json 2> $O/bench_${TAG}_${N}gpu.err; echo "bench rc=$?"; tail -c 2500 $O/bench_${TAG}_${N}gpu.json; grep -c "NCCL INFO" $O/bench_${TAG}_${N}gpu.err; grep -m3 "nranks\|NVLS\|Connected all" $O/bench_${TAG}_${N}gpu.err | cut -c1-200
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > $O/bench_ref_${TAG}_${N}gpu.json 2> /dev/null; echo "ref rc=$?"; tail -c 700 $O/bench_ref_${TAG}_${N}gpu.json ;;
d2h)
  N=$(nvidia-smi -L | wc -l)
  { nvidia-smi topo -m 2>&1 | head -24; lscpu 2>/dev/null | grep -i "numa\|socket\|^CPU(s)\|model name"; echo "cgroup cpus: $(nproc)"
    NCCL_DEBUG=WARN timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 tools/d2h_probe.py 2>&1 | grep "^{"
  } > $O/d2h_probe_${TAG}_${N}gpu.txt 2>&1; echo "d2h rc=$?"; cat $O/d2h_probe_${TAG}_${N}gpu.txt ;;
multicfg)
  N=$(nvidia-smi -L | wc -l)
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 tools/run_configs.py --configs 3,5 --c4-chains 4194304 --out $O/configs_${TAG}_${N}gpu.jsonl > $O/configs_${TAG}_${N}gpu.log 2>&1; echo "configs rc=$?"; grep "^{" $O/configs_${TAG}_${N}gpu.log | cut -c1-400 ;;
drawab)
  timeout 300 python -m pytest tests/test_fullsize_slices_gpu.py tests/test_rng_gpu.py tests/test_parity_gpu.py -m gpu -q --timeout 120 -k "config2 or config4 or rng or native or hmc_parity or rwmh" -p no:cacheprovider > $O/pytest_draw_$TAG.log 2>&1; echo "draw tests rc=$?"; tail -3 $O/pytest_draw_$TAG.log
  { for v in _b0 _b2 "" _b10; do lib=$PWD/mcmc_b200/csrc/libmcmc_b200$v.so; [ -f $lib ] || continue; echo "== draw batch variant '$v'"
      MCMC_B200_LIB=$lib timeout 120 python tools/perf_hmc.py --samples 64 --Ls 16,32,64,128 --eps 0.02
      MCMC_B200_LIB=$lib timeout 120 python tools/perf_hmc.py --sampler rwmh --target standard_normal --chains 4194304 --samples 64 --Ls 1 --eps 0.532; done
  } > $O/drawab_$TAG.txt 2>&1; cat $O/drawab_$TAG.txt ;;
kinab)
  { for v in "" _k4 _e1 _k4e1; do lib=$PWD/mcmc_b200/csrc/libmcmc_b200$v.so; [ -f $lib ] || continue; echo "== hmc variant '$v' (k = kinetic sum ways, e = early state load)"
      MCMC_B200_LIB=$lib timeout 120 python tools/perf_hmc.py --samples 64 --Ls 16,32,64,128 --eps 0.02
      MCMC_B200_LIB=$lib timeout 120 python tools/perf_hmc.py --target standard_normal --samples 64 --Ls 16 --eps 0.3
      MCMC_B200_LIB=$lib timeout 120 python tools/perf_hmc.py --target dense_gaussian --samples 64 --Ls 16 --eps 0.1; done
  } > $O/kinab_$TAG.txt 2>&1; cat $O/kinab_$TAG.txt ;;
nutschunk)
  { for ch in 0 1 2 4 8; do for t in standard_normal rosenbrock neals_funnel; do e=0.5; [ $t = rosenbrock ] && e=0.02; [ $t = neals_funnel ] && e=0.1
      echo "MCMC_NUTS_CHUNK=$ch"; MCMC_NUTS_CHUNK=$ch timeout 300 python tools/perf_hmc.py --sampler nuts --target $t --samples 96 --Ls 1 --eps $e; done; done
  } > $O/nuts_chunk_$TAG.txt 2>&1; cat $O/nuts_chunk_$TAG.txt ;;
slices)
  timeout 900 python -m pytest tests/test_fullsize_slices_gpu.py tests/test_grid_search_gpu.py tests/test_diagnostics_gpu.py tests/test_quality_gpu.py -m gpu -q -p no:cacheprovider > $O/pytest_slices_$TAG.log 2>&1; echo "slices rc=$?"; tail -8 $O/pytest_slices_$TAG.log ;;
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
