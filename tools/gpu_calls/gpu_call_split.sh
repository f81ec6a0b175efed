#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_cuda_parity.py -m gpu -x -q -k wavefront > gpurun_out/sp_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/sp_pytest.log
rm -f gpurun_out/sp_ab.txt
bash tools/ab_env.sh gpurun_out/sp_ab.txt "RAYCA_WAVE_SPLIT=0" "RAYCA_WAVE_FETCH=0" "RAYCA_WAVE_FETCH_MIN=12" "RAYCA_WAVE_FETCH_MIN=16" "RAYCA_WAVE_FETCH_MIN=20" "RAYCA_WAVE_FETCH_MIN=24" "RAYCA_WAVE_FETCH_MIN=28" "RAYCA_WAVE_FETCH_MIN=32" -- --steps 20 --warmup 3 --max-depth 2
bash tools/ab_env.sh gpurun_out/sp_ab.txt "RAYCA_WAVE_SPLIT=0" "RAYCA_WAVE_FETCH=0" "RAYCA_WAVE_FETCH=1" -- --steps 10 --warmup 3 --max-depth 4
ncu -k regex:"wave_|render_kernel|RadixSort" --metrics gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,sm__warps_active.avg.pct_of_peak_sustained_active --clock-control none -c 60 --csv --log-file gpurun_out/sp_launches.csv python bench.py --no-cpu-baseline --steps 2 --warmup 1 --max-depth 2 > gpurun_out/sp_ncu.log 2>&1
tail -3 gpurun_out/sp_pytest.log; grep -E "^==|^value" gpurun_out/sp_ab.txt
