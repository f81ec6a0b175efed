#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_cuda_parity.py tests/test_cuda_multi_device.py tests/test_cuda_edge_cases.py -m gpu -x -q -k "shard or frame or multi or deep or ggx or progressive" > gpurun_out/r02_gpu_tests7.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_gpu_tests7.log
tail -4 gpurun_out/r02_gpu_tests7.log | cut -c1-250
out=gpurun_out/r02_shard_ab2.txt; rm -f $out
python tools/shard_ab.py --shards 8 >> $out 2>&1
python tools/shard_ab.py --shards 8 --flags 8 >> $out 2>&1
python tools/shard_ab.py --shards 4 --flags 8 >> $out 2>&1
python tools/shard_ab.py --shards 2 --flags 8 >> $out 2>&1
python tools/shard_ab.py --shards 8 --flags 8 --max-depth 2 >> $out 2>&1
grep "^lib" $out | cut -c1-400
