#!/bin/bash
# final check of the round: tests, smoke, the default bench line, the bounce line, the soup's scene_create phases
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/final_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/final_pytest.log
tail -4 gpurun_out/final_pytest.log | cut -c1-200
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; echo "bench rc=$?"
timeout 900 python bench.py --steps 20 --warmup 3 --max-depth 2 > gpurun_out/final_bounce.json 2> gpurun_out/final_bounce.err; echo "bounce rc=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 2>/dev/null | tail -1 | cut -c1-200
python tools/draw_phases.py --workload soup --width 1920 --height 1080 > gpurun_out/final_phases_soup.txt 2>&1; tail -24 gpurun_out/final_phases_soup.txt
python - <<'PY'
import json
for n in ('bench','bounce'):
    d=json.loads(open(f'gpurun_out/final_{n}.json').read().strip().splitlines()[-1])
    print(n,'value',round(d['value'],1),'e2e',round(d['e2e']['value'],1),'launches',d['gpu_launches'],'parity_ok',d.get('parity_ok'),'crc',d.get('frame_crc32'),'draw',d.get('draw_e2e',{}).get('library_ms'),'frac',round(d['roofline']['frac'],3),'traffic',d['roofline']['traffic'])
PY
