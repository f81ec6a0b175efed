#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_gpu_tests12.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_gpu_tests12.log
tail -12 gpurun_out/r02_gpu_tests12.log | cut -c1-250
out=gpurun_out/r02_wave_ab.txt; rm -f $out
run() { echo "== $*" >> $out; env "$@" python bench.py --no-cpu-baseline --steps 20 --max-depth 2 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print('value', round(d['value'],1), 'kernel_ms', round(d['kernel_ms'],4), 'e2e', round(d['e2e']['value'],1), 'crc', d.get('frame_crc32'), 'launches', d.get('gpu_launches'))
    elif 'Error' in l or 'error' in l: print(l[:300])
" >> $out; }
run RAYCA_NO_WAVEFRONT=1
run RAYCA_WAVE_KEY_BITS=7
run RAYCA_WAVE_KEY_BITS=5
run RAYCA_WAVE_KEY_BITS=6
run RAYCA_WAVE_KEY_BITS=8
run RAYCA_WAVE_KEY_BITS=9
cat $out
