#!/bin/bash
# GPU call 3 (round 2): A/B of the walk / shading variants on the three workloads.
mkdir -p gpurun_out
rm -f gpurun_out/r02_ab1.txt
timeout 1200 tools/ab2.sh gpurun_out/r02_ab1.txt base sm8 sm12 sm16 fold fold_sm12 sh all12 soa_sm12 soa_sm12_sh -- --steps 40
timeout 600 tools/ab2.sh gpurun_out/r02_ab1.txt base sm12 all12 soa_sm12_sh -- --steps 20 --max-depth 2
timeout 900 tools/ab2.sh gpurun_out/r02_ab1.txt base sm12 all12 soa_sm12_sh -- --steps 3 --workload soup --width 3840 --height 2160 --spp 16
grep -E "^==|^value" gpurun_out/r02_ab1.txt | cut -c1-160
