#!/bin/bash
# usage: tools/gpu_scale.sh N   — run under `gpurun --gpus N`: tests on N real devices, the one-call multi-device context, and the
# torchrun bench lines of the three workloads at N GPUs.
N=$1
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r02_scale_n${N}_smi.txt 2>&1
timeout 600 python -m pytest tests/test_cuda_multi_device.py -m gpu -x -q > gpurun_out/r02_scale_n${N}_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_scale_n${N}_tests.log
tail -3 gpurun_out/r02_scale_n${N}_tests.log | cut -c1-250
timeout 300 python tools/multi_device_bench.py > gpurun_out/r02_multi_device_n${N}.json 2> gpurun_out/r02_multi_device_n${N}.err; echo "multi rc=$?"
timeout 300 python tools/multi_device_bench.py --max-depth 2 > gpurun_out/r02_multi_device_bounce_n${N}.json 2>> gpurun_out/r02_multi_device_n${N}.err; echo "multi bounce rc=$?"
run() { # name, args
  local name=$1; shift
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@" > gpurun_out/r02_scale_${name}_n${N}.json 2> gpurun_out/r02_scale_${name}_n${N}.err; echo "$name rc=$?"
}
run atrium --steps 20 --warmup 3
run bounce --steps 20 --warmup 3 --max-depth 2
run soup --steps 3 --warmup 3 --workload soup --width 3840 --height 2160 --spp 16
for f in gpurun_out/r02_scale_*_n${N}.json gpurun_out/r02_multi_device*_n${N}.json; do echo "== $f"; tail -1 $f | cut -c1-700; done
