#!/bin/bash
# same-box A/B of NUTS kernel variants: a = HEAD, b = sign-folded momentum + eps*w drift, c = b + kinetic energy from the eps*w table
O=gpurun_out; mkdir -p $O
{ for rep in 1 2 3; do for v in _a _b _c; do lib=$PWD/mcmc_b200/csrc/libmcmc_b200$v.so
  for t in standard_normal rosenbrock neals_funnel; do e=0.5; [ $t = rosenbrock ] && e=0.02; [ $t = neals_funnel ] && e=0.1
    echo -n "rep=$rep variant=$v "; MCMC_B200_LIB=$lib timeout 100 python tools/perf_hmc.py --sampler nuts --target $t --samples 96 --Ls 1 --eps $e; done; done; done
} > $O/nuts_abc.txt 2>&1
python - <<'PY'
import json,re,collections
d=collections.defaultdict(list)
for line in open('gpurun_out/nuts_abc.txt'):
    m=re.match(r'rep=(\d) variant=(\S+) (\{.*\})',line)
    if m:
        j=json.loads(m.group(3)); d[(m.group(2),j['target'])].append(j['ms'])
for k in sorted(d): print(k, ' '.join(f'{x:.3f}' for x in d[k]), ' min %.3f'%min(d[k]))
PY
