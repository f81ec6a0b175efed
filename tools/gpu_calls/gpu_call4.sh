#!/bin/bash
# GPU call 4 (round 2): new tests (multi-device context, frame time-out), ncu of the bounce and soup kernels, launch list.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_gpu_tests4.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_gpu_tests4.log
tail -15 gpurun_out/r02_gpu_tests4.log | cut -c1-300
timeout 600 python bench.py --steps 20 --warmup 3 --max-depth 2 > gpurun_out/r02_bench_bounce.json 2> gpurun_out/r02_bench_bounce.err; echo "bounce rc=$?"
timeout 900 python bench.py --steps 3 --warmup 3 --workload soup --width 3840 --height 2160 --spp 16 --cpu-rows 2 > gpurun_out/r02_bench_soup.json 2> gpurun_out/r02_bench_soup.err; echo "soup rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 4 -c 1 -o gpurun_out/r02_bounce python bench.py --steps 2 --warmup 3 --no-cpu-baseline --max-depth 2 > gpurun_out/r02_ncu_bounce.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 4 -c 1 -o gpurun_out/r02_soup python bench.py --steps 2 --warmup 3 --no-cpu-baseline --workload soup --width 3840 --height 2160 --spp 16 > gpurun_out/r02_ncu_soup.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r02_launches_atrium.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r02_launches.log 2>&1
ls -la gpurun_out | grep r02_ | cut -c30-
