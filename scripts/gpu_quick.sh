#!/bin/bash
# quick GPU visit: parity tests + default bench (no CPU baseline)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "rc=$?" >> gpurun_out/bench_default.err
tail -3 gpurun_out/pytest_gpu.log; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_default.json').read().strip().splitlines()[-1])
print(d['value'], d['e2e']['value'], d['roofline'].get('kernels_ms'), d['roofline']['frac'], d['roofline']['whole_step_frac'])
PY
