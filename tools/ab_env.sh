#!/bin/bash
# usage: tools/ab_env.sh <outfile> "<ENV=VAL ...>" ... -- <bench args> ; one bench.py run per environment string
out=$1; shift
envs=()
while [ "$1" != "--" ]; do envs+=("$1"); shift; done
shift
for e in "${envs[@]}"; do
  echo "== $e" >> $out
  env $e python bench.py --no-cpu-baseline "$@" 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print('value', round(d['value'],1), 'kernel_ms', round(d['kernel_ms'],4), 'e2e', round(d['e2e']['value'],1), 'nodes/ray', round(d['walk']['node_records_per_ray'],2), 'tri B/ray', round(d['walk']['triangle_bytes_per_ray'],1), 'build_s', round(d['build_s'],2))
    else: print(l)
" >> $out
done
