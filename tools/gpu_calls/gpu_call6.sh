#!/bin/bash
# GPU call 6: per-rank kernel time of an 8-way (and 4-, 2-way) split, emulated on one GPU (tools/shard_ab.py).
mkdir -p gpurun_out
out=gpurun_out/r02_shard_ab.txt; rm -f $out
python tools/shard_ab.py --shards 8 >> $out 2>&1
python tools/shard_ab.py --shards 8 --flags 2 >> $out 2>&1   # persistent scheduler
python tools/shard_ab.py --shards 8 --tile 16 >> $out 2>&1
python tools/shard_ab.py --shards 8 --tile 32 >> $out 2>&1
for v in w21 w11 w42 mb9; do RAYCA_CUDA_LIB=rayca_b200/variants/$v.so python tools/shard_ab.py --shards 8 >> $out 2>&1; done
python tools/shard_ab.py --shards 4 >> $out 2>&1
python tools/shard_ab.py --shards 2 >> $out 2>&1
python tools/shard_ab.py --shards 8 --max-depth 2 >> $out 2>&1
grep "^lib" $out | cut -c1-400
