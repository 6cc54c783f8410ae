#!/bin/bash
# 8-GPU visit: the driver's own launch line for N=8 (and N=4, N=2), no CPU baseline
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/scale_gpus.txt 2>&1
for n in 8 4 2; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2955$n bench.py --gpus $n --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; echo "rc=$?" >> gpurun_out/bench_n$n.err
done
timeout 600 python bench.py --gpus 1 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err
for n in 1 2 4 8; do python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/bench_n$n.json') if l.startswith('{')][-1])
    print($n, d['value'], d['e2e']['value'], d['ms_per_step'], d.get('clocks'))
except Exception as e:
    print($n, 'ERR', e); print(open('gpurun_out/bench_n$n.err').read()[-1500:])
PY
done
