#!/bin/bash
# final 8-GPU visit: multi-GPU tests, then the driver's own launch line for N = 8 and N = 2 (no CPU baseline)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "sharded or device_list or multi" > gpurun_out/pytest_multi.log 2>&1; tail -3 gpurun_out/pytest_multi.log
for n in 8 2; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2955$n bench.py --gpus $n --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; echo "rc=$?" >> gpurun_out/bench_n$n.err
  python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/bench_n$n.json') if l.startswith('{')][-1])
    sp=d.get('sharded_product') or {}
    print($n, d['value'], d['e2e']['value'], d['ms_per_step'], sp.get('value'), sp.get('is_one'), (d.get('single_call_multi_gpu') or {}).get('value'))
except Exception as e:
    print($n, 'ERR', e); print(open('gpurun_out/bench_n$n.err').read()[-1500:])
PY
done
