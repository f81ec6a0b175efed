#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/fs_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/fs_pytest.log
tail -4 gpurun_out/fs_pytest.log | cut -c1-250
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/fs_smoke.log 2>&1; tail -1 gpurun_out/fs_smoke.log
for k in wave_walk_fetch_kernel wave_shade_kernel; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 4 -c 1 -o /tmp/$k python bench.py --steps 2 --warmup 3 --no-cpu-baseline --max-depth 2 > gpurun_out/fs_ncu_$k.log 2>&1
  python tools/ncu_summary.py /tmp/$k.ncu-rep gpurun_out/fs_ncu_$k.json "ncu --set full --clock-control none --import-source on -k regex:$k -s 4 -c 1 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --max-depth 2" > /dev/null 2>&1
  rm -f /tmp/$k.ncu-rep
done
timeout 900 python bench.py --steps 20 --warmup 3 --max-depth 2 > gpurun_out/fs_bench_bounce.json 2> gpurun_out/fs_bench_bounce.err; echo "bounce rc=$?"
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/fs_bench_default.json 2> gpurun_out/fs_bench_default.err; echo "default rc=$?"
rm -f gpurun_out/fs_thresh.txt
for h in 540 360 270; do
  bash tools/ab_env.sh gpurun_out/fs_thresh.txt "RAYCA_NO_WAVEFRONT=1 H=$h" "RAYCA_WAVE_MIN_PATHS=0 H=$h" -- --steps 20 --warmup 3 --max-depth 2 --height $h
done
grep -E "^==|^value" gpurun_out/fs_thresh.txt | sed 's/ nodes\/ray.*//'
python - <<'PY'
import json
for n in ('bounce','default'):
    d=json.loads(open(f'gpurun_out/fs_bench_{n}.json').read().strip().splitlines()[-1])
    print(n,'value',round(d['value'],1),'kernel_ms',round(d['kernel_ms'],4),'e2e',round(d['e2e']['value'],1),'launches',d['gpu_launches'],'parity_ok',d.get('parity_ok'),'crc',d.get('frame_crc32'))
for k in ('wave_walk_fetch_kernel','wave_shade_kernel'):
    d=json.load(open(f'gpurun_out/fs_ncu_{k}.json'))['launches'][0]
    print(k,{x:d.get(x) for x in ('gpu__time_duration.sum','smsp__thread_inst_executed_per_inst_executed.ratio','smsp__issue_active.avg.pct_of_peak_sustained_active','lts__t_sector_hit_rate.pct','l1tex__t_sector_hit_rate.pct','smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio','sm__warps_active.avg.pct_of_peak_sustained_active','dram__bytes_read.sum','dram__bytes_write.sum','launch__registers_per_thread')})
PY
