#!/bin/bash
# GPU call 5 (round 2): tests, bounce / soup / flighthelmet bench lines, ncu summaries made ON THE BOX (reports deleted: the
# merge back is limited to 64 MiB), launch list, launch-shape A/B, Draw::draw phases.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_gpu_tests5.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_gpu_tests5.log
tail -6 gpurun_out/r02_gpu_tests5.log | cut -c1-250
timeout 600 python bench.py --steps 20 --warmup 3 --max-depth 2 > gpurun_out/r02_bench_bounce.json 2> gpurun_out/r02_bench_bounce.err; echo "bounce rc=$?"
timeout 600 python bench.py --steps 20 --warmup 3 --workload flighthelmet > gpurun_out/r02_bench_flighthelmet.json 2> gpurun_out/r02_bench_flighthelmet.err; echo "flighthelmet rc=$?"
timeout 900 python bench.py --steps 3 --warmup 3 --workload soup --width 3840 --height 2160 --spp 16 > gpurun_out/r02_bench_soup.json 2> gpurun_out/r02_bench_soup.err; echo "soup rc=$?"
ncu_one() { # name, bench args...
  local name=$1; shift
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 4 -c 1 -o /tmp/$name python bench.py --steps 2 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/r02_ncu_$name.log 2>&1
  python tools/ncu_summary.py /tmp/$name.ncu-rep gpurun_out/r02_ncu_$name.json "ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 4 -c 1 python bench.py --steps 2 --warmup 3 --no-cpu-baseline $*" > /dev/null 2>&1
  ncu -i /tmp/$name.ncu-rep --page source --csv > gpurun_out/r02_source_$name.csv 2>/dev/null
  rm -f /tmp/$name.ncu-rep
}
ncu_one atrium
ncu_one bounce --max-depth 2
ncu_one flighthelmet --workload flighthelmet
ncu_one soup --workload soup --width 3840 --height 2160 --spp 16
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r02_launches_atrium.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r02_launches.log 2>&1
rm -f gpurun_out/r02_ab2.txt
timeout 1200 tools/ab2.sh gpurun_out/r02_ab2.txt base mb9 mb10 mb12 w21 w11 w42 w12 w41 -- --steps 40
timeout 600 tools/ab2.sh gpurun_out/r02_ab2.txt base mb10 w21 w11 -- --steps 20 --max-depth 2
RAYCA_TIMING=1 timeout 300 python tools/draw_phases.py > gpurun_out/r02_draw_phases.txt 2>&1
grep -E "^==|^value" gpurun_out/r02_ab2.txt | sed 's/ clk .*//' | cut -c1-170
du -sh gpurun_out
