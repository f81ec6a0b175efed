#!/bin/bash
# GPU call 1 (round 2): parity tests, A/B of the walk / shading variants, the new bench line, one ncu capture.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv > gpurun_out/r02_smi.txt 2>&1
nproc >> gpurun_out/r02_smi.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_gpu_tests.log
tail -5 gpurun_out/r02_gpu_tests.log
rm -f gpurun_out/r02_ab1.txt
timeout 900 tools/ab2.sh gpurun_out/r02_ab1.txt base sm8 sm12 sm16 fold fold_sm12 sh all12 soa_sm12 soa_sm12_sh -- --steps 40
timeout 600 tools/ab2.sh gpurun_out/r02_ab1.txt base sm12 all12 soa_sm12_sh -- --steps 20 --max-depth 2
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r02_bench_default.json 2> gpurun_out/r02_bench_default.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/r02_bench_default.json
timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 4 -c 1 -o gpurun_out/r02_atrium_base python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r02_ncu_base.log 2>&1
ls -la gpurun_out | tail -20
