#!/bin/bash
# A/B of NUTS builds at config-5 shape: tools/nuts_ab.sh LIB...   (LIB = path of a build of the same ABI)
for lib in "$@"; do
  for spec in "standard_normal 0.5" "rosenbrock 0.02" "neals_funnel 0.1"; do
    set -- $spec
    MCMC_B200_LIB=$PWD/$lib python tools/perf_hmc.py --sampler nuts --target $1 --samples 96 --Ls 1 --eps $2 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try:
        d=json.loads(l); print('$lib', d['target'], 'ms', round(d['ms'],3), 'G grad-evals/s', round(d['grad_evals_per_s']/1e9,3))
    except Exception: print(l.strip())
"
  done
done
