#!/bin/bash
# parity tests, then large batches with and without the 512-thread blocks (device-resident, bench.py --batch-log2)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -2 gpurun_out/pytest_gpu.log
for k in 16 17 18 19 20; do
  for big in 1 0; do
    BN254_BIG_BLOCKS=$big timeout 600 python bench.py --batch-log2 $k --steps 2 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('2^$k big_blocks=$big', round(d['value']), round(d['e2e']['value']), round(d['ms_per_step'],3), d['clocks']['reasons'])"
  done
done | tee gpurun_out/big_blocks.txt
