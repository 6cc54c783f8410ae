#!/bin/bash
# batch-size sweep of the default path against the local-memory kernels (device-resident, bench.py --batch-log2)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -2 gpurun_out/pytest_gpu.log
for k in 12 13 14 15 16 17 18; do
  for mode in new old; do
    if [ $mode = old ]; then export BN254_FEXP_SMEM=0 BN254_MILLER_SMEM=0; else unset BN254_FEXP_SMEM BN254_MILLER_SMEM; fi
    timeout 600 python bench.py --batch-log2 $k --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('2^$k $mode', round(d['value']), round(d['e2e']['value']), round(d['ms_per_step'],3))"
  done
done | tee gpurun_out/sweep.txt
