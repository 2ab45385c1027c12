#!/bin/bash
# A/B of RWMH kernel builds (config 4 shape scaled to 1M and 4M chains): tools/rwmh_sweep.sh SUFFIX...
for v in "$@"; do
  lib=mcmc_b200/csrc/libmcmc_b200_$v.so; env=""
  [ "$v" = unrolled ] && { lib=mcmc_b200/csrc/libmcmc_b200_rwu5.so; export MCMC_RWMH_UNROLLED=1; } || unset MCMC_RWMH_UNROLLED
  for C in 1048576 4194304; do
    MCMC_B200_LIB=$PWD/$lib python tools/perf_hmc.py --sampler rwmh --target standard_normal --chains $C --samples 64 --Ls 1 --eps 0.532 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try:
        d=json.loads(l); print('$v', d['chains'], 'ms', round(d['ms'],3), 'TB/s', round(d['chains']*d['samples']*168/d['ms']/1e9,3), 'accept', round(d['accept'],4))
    except Exception: print(l.strip())
"
  done
done
