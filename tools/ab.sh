#!/bin/bash
# usage: tools/ab.sh <outfile> <bench args...> ; runs bench.py for every variant .so and both grid modes
out=$1; shift
for so in rayca_b200/variants/*.so; do
  for fl in 0 2; do
    echo "== $(basename $so) flags=$fl" >> $out
    RAYCA_CUDA_LIB=$so python bench.py --no-cpu-baseline --flags $fl "$@" 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print('value', round(d['value'],1), 'kernel_ms', round(d['kernel_ms'],4), 'e2e', round(d['e2e']['value'],1), 'frac', round(d['roofline']['frac'],3), 'crc', d.get('frame_crc32'))
    else: print(l)
" >> $out
  done
done
