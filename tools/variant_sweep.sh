#!/bin/bash
# A/B of experimental builds of the same ABI (MCMC_B200_LIB): tools/variant_sweep.sh SUFFIX...
for v in "$@"; do
  echo "== variant $v"
  lib=mcmc_b200/csrc/libmcmc_b200_$v.so
  [ "$v" = base ] && lib=mcmc_b200/csrc/libmcmc_b200.so
  MCMC_B200_LIB=$PWD/$lib python tools/perf_hmc.py --samples 64 --Ls 16,32,64,128 --eps 0.02 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l); print(d['L'], round(d['ms'],4), round(d['grad_evals_per_s']/1e9,2), round(d['accept'],4))
    except Exception: print(l.strip())
"
done
