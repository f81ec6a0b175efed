#!/bin/bash
N=$1
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 --max-depth 2 > gpurun_out/r02_scale_bounce_wave_n${N}.json 2> gpurun_out/r02_scale_bounce_wave_n${N}.err; echo "bounce rc=$?"
tail -1 gpurun_out/r02_scale_bounce_wave_n${N}.json | cut -c1-260
