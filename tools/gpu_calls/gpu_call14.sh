#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_gpu_tests14.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_gpu_tests14.log
tail -6 gpurun_out/r02_gpu_tests14.log | cut -c1-250
timeout 600 python bench.py --steps 20 --warmup 3 --max-depth 2 > gpurun_out/r02_bench_bounce_wave.json 2> gpurun_out/r02_bench_bounce_wave.err; echo "bounce rc=$?"
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r02_bench_default2.json 2> gpurun_out/r02_bench_default2.err; echo "default rc=$?"
for k in wave_stage_kernel wave_stage0_kernel; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 4 -c 1 -o /tmp/$k python bench.py --steps 2 --warmup 3 --no-cpu-baseline --max-depth 2 > gpurun_out/r02_ncu_$k.log 2>&1
  python tools/ncu_summary.py /tmp/$k.ncu-rep gpurun_out/r02_ncu_$k.json "ncu --set full --clock-control none --import-source on -k regex:$k -s 4 -c 1 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --max-depth 2" > /dev/null 2>&1
  rm -f /tmp/$k.ncu-rep
done
python - <<'PY'
import json
for n in ('bounce_wave','default2'):
    d=json.loads(open(f'gpurun_out/r02_bench_{n}.json').read().strip().splitlines()[-1])
    print(n,'value',round(d['value'],1),'kernel_ms',round(d['kernel_ms'],4),'e2e',round(d['e2e']['value'],1),'launches',d['gpu_launches'],'parity_ok',d.get('parity_ok'),'draw',d.get('draw_e2e'))
    print('  rows',{k:v for k,v in d['parity_rows'].items() if k not in('rows','against')}); print('  frame',{k:v for k,v in d['parity_frame'].items() if k!='against'})
for k in ('wave_stage_kernel','wave_stage0_kernel'):
    d=json.load(open(f'gpurun_out/r02_ncu_{k}.json'))['launches'][0]
    print(k,{x:d[x] for x in ('gpu__time_duration.sum','smsp__thread_inst_executed_per_inst_executed.ratio','smsp__issue_active.avg.pct_of_peak_sustained_active','lts__t_sector_hit_rate.pct','l1tex__t_sector_hit_rate.pct','smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio','sm__warps_active.avg.pct_of_peak_sustained_active','dram__bytes_read.sum','dram__bytes_write.sum')})
PY
