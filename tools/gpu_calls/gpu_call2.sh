#!/bin/bash
# GPU call 2 (round 2): the allocation-free entry point (alone, then after every other GPU test), then the A/B runs.
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r02_no_alloc_alone.log 2>&1; echo "alone rc=$?" >> gpurun_out/r02_no_alloc_alone.log
tail -4 gpurun_out/r02_no_alloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_gpu_tests2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_gpu_tests2.log
tail -4 gpurun_out/r02_gpu_tests2.log
rm -f gpurun_out/r02_ab1.txt
timeout 1200 tools/ab2.sh gpurun_out/r02_ab1.txt base sm8 sm12 sm16 fold fold_sm12 sh all12 soa_sm12 soa_sm12_sh -- --steps 40
timeout 600 tools/ab2.sh gpurun_out/r02_ab1.txt base sm12 all12 soa_sm12_sh -- --steps 20 --max-depth 2
timeout 900 tools/ab2.sh gpurun_out/r02_ab1.txt base sm12 all12 soa_sm12_sh -- --steps 3 --workload soup --width 3840 --height 2160 --spp 16
cat gpurun_out/r02_ab1.txt
