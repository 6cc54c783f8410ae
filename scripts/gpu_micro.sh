#!/bin/bash
mkdir -p gpurun_out
./tools/microbench/pipes.bin 1965 > gpurun_out/pipes.txt 2>&1
for k in 16 18 20; do
  for v in "" "_r255"; do
    timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --batch-log2 $k --variant "$v" > gpurun_out/bench_k${k}${v}.json 2> gpurun_out/bench_k${k}${v}.err
  done
done
cat gpurun_out/pipes.txt
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/bench_k*.json")):
    try:
        d=json.load(open(f)); r=d['roofline']; print(f, "value %.0f"%d['value'], "e2e %.0f"%d['e2e']['value'], r['kernels_ms'])
    except Exception as e: print(f, "FAILED", e)
PY
