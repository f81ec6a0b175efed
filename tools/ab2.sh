#!/bin/bash
# usage: tools/ab2.sh <outfile> <variant names...> -- <bench args...> ; static grid only, prints the frame crc
out=$1; shift
names=()
while [ "$1" != "--" ] && [ $# -gt 0 ]; do names+=("$1"); shift; done
shift
for n in "${names[@]}"; do
  echo "== $n $*" >> $out
  RAYCA_AB_FRAME=/tmp/ab_frame_$(echo "$*" | tr -c "a-zA-Z0-9" _).npy RAYCA_CUDA_LIB=rayca_b200/variants/$n.so python bench.py --no-cpu-baseline "$@" 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print('value', round(d['value'],1), 'kernel_ms', round(d['kernel_ms'],4), 'e2e', round(d['e2e']['value'],1), 'l2frac', round(d['roofline']['frac'],3), 'crc', d.get('frame_crc32'), 'clk', d['clocks'])
    else: print(l)
" >> $out
done
