#!/bin/bash
# same-box A/B of NUTS kernel variants (library suffixes as arguments; default: _a _b _c of the sign-folding experiment,
# a = round-2 kernel, b = sign-folded momentum + eps*w drift, c = b + kinetic energy from the eps*w table)
VARIANTS=${*:-"_a _b _c"}
O=gpurun_out; mkdir -p $O
{ for rep in 1 2 3; do for v in $VARIANTS; do lib=$PWD/mcmc_b200/csrc/libmcmc_b200$v.so
  for t in standard_normal rosenbrock neals_funnel; do e=0.5; [ $t = rosenbrock ] && e=0.02; [ $t = neals_funnel ] && e=0.1
    echo -n "rep=$rep variant=$v "; MCMC_B200_LIB=$lib timeout 100 python tools/perf_hmc.py --sampler nuts --target $t --samples 96 --Ls 1 --eps $e; done; done; done
  for v in $VARIANTS; do echo "== latency probe $v"; MCMC_B200_LIB=$PWD/mcmc_b200/csrc/libmcmc_b200$v.so MCMC_NUTS_THREAD=1 timeout 150 python tools/nuts_latency_probe.py 2>&1 | grep -v "^==" | grep "C=    1\|C=   32"; done
} > $O/nuts_abc.txt 2>&1
python - <<'PY'
import json,re,collections
d=collections.defaultdict(list)
for line in open('gpurun_out/nuts_abc.txt'):
    m=re.match(r'rep=(\d) variant=(\S+) (\{.*\})',line)
    if m:
        j=json.loads(m.group(3)); d[(m.group(2),j['target'])].append(j['ms'])
for k in sorted(d): print(k, ' '.join(f'{x:.3f}' for x in d[k]), ' min %.3f'%min(d[k]))
print(''.join(l for l in open('gpurun_out/nuts_abc.txt') if 'latency probe' in l or 'per step' in l))
PY
