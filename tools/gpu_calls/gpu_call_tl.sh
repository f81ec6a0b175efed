#!/bin/bash
# sanity (tests, smoke, default bench) on the shipped build, then the shared-memory treelet A/B
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/tl_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/tl_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/tl_smoke.log 2>&1
python bench.py --steps 20 --warmup 3 > gpurun_out/tl_bench.log 2>&1
rm -f gpurun_out/tl_ab.txt
for rep in 1 2; do
  bash tools/ab2.sh gpurun_out/tl_ab.txt base tl64 tl256 -- --steps 30 --warmup 5
done
bash tools/ab2.sh gpurun_out/tl_ab.txt base tl64 tl256 -- --steps 10 --warmup 3 --max-depth 2 --flags 16
tail -3 gpurun_out/tl_pytest.log; tail -2 gpurun_out/tl_smoke.log; tail -1 gpurun_out/tl_bench.log | cut -c1-400; grep -E "^==|^value" gpurun_out/tl_ab.txt
