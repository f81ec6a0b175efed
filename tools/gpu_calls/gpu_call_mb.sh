#!/bin/bash
mkdir -p gpurun_out
python tools/shard_ab.py --shards 8 --reps 30 >> gpurun_out/warm_shard.txt 2>&1
python tools/shard_ab.py --shards 8 --reps 30 --no-flush >> gpurun_out/warm_shard.txt 2>&1
python tools/shard_ab.py --shards 4 --reps 30 --no-flush >> gpurun_out/warm_shard.txt 2>&1
grep "lib" gpurun_out/warm_shard.txt
