#!/bin/bash
mkdir -p gpurun_out
out=gpurun_out/r02_wave_ab2.txt; rm -f $out
run() { echo "== $*" >> $out; env "$@" python bench.py --no-cpu-baseline --steps 20 --max-depth 2 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print('value', round(d['value'],1), 'kernel_ms', round(d['kernel_ms'],4), 'e2e', round(d['e2e']['value'],1), 'crc', d.get('frame_crc32'))
    elif 'Error' in l or 'error' in l: print(l[:300])
" >> $out; }
run RAYCA_NO_WAVEFRONT=1
run RAYCA_WAVE_KEY_MODE=0 RAYCA_WAVE_KEY_BITS=6
for kb in 3 4 5; do for db in 3 4 5 6; do run RAYCA_WAVE_KEY_MODE=1 RAYCA_WAVE_KEY_BITS=$kb RAYCA_WAVE_DIR_BITS=$db; done; done
run RAYCA_WAVE_KEY_MODE=2 RAYCA_WAVE_KEY_BITS=4 RAYCA_WAVE_DIR_BITS=4
run RAYCA_WAVE_KEY_MODE=2 RAYCA_WAVE_KEY_BITS=5 RAYCA_WAVE_DIR_BITS=3
cat $out
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"wave|Radix|render_kernel" -s 60 -c 14 --csv --log-file gpurun_out/r02_wave_launches.csv python bench.py --no-cpu-baseline --steps 4 --max-depth 2 > /dev/null 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(l for l in open('gpurun_out/r02_wave_launches.csv') if l.startswith('"'))]
h=rows[0]; k=h.index('Kernel Name'); v=h.index('Metric Value')
for r in rows[1:]: print(r[k][:70], r[v])
PY
