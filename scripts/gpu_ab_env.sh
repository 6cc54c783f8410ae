#!/bin/bash
# A/B of one environment switch: parity tests + default bench with and without it.  usage: gpu_ab_env.sh VAR=VALUE
mkdir -p gpurun_out
export "$1"
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_ab.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_ab.log
tail -3 gpurun_out/pytest_gpu_ab.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_ab_on.json 2> gpurun_out/bench_ab_on.err
unset "${1%%=*}"
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_ab_off.json 2> gpurun_out/bench_ab_off.err
python - <<'PY'
import json
for t in ('on','off'):
    try:
        d=json.loads(open('gpurun_out/bench_ab_%s.json'%t).read().strip().splitlines()[-1])
        print(t, round(d['value']), round(d['e2e']['value']), {k: round(v,3) for k,v in d['roofline'].get('kernels_ms',{}).items()})
    except Exception as e:
        print(t, 'failed', e, open('gpurun_out/bench_ab_%s.err'%t).read()[-600:])
PY
