#!/bin/bash
# usage: gpu_variants.sh "<variant list>" "<batch-log2 list>"  -- bench each library variant (device-resident value only)
mkdir -p gpurun_out
for k in ${2:-16}; do
for v in $1; do
  vv=$v; [ "$v" = "default" ] && vv=""
  timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --batch-log2 $k --variant "$vv" > gpurun_out/bench_var_${v}_k$k.json 2> gpurun_out/bench_var_${v}_k$k.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/bench_var_${v}_k$k.json")); r=d['roofline']
    print("$v k=$k", "value %.0f"%d['value'], "ms/step %.2f"%d['ms_per_step'], "e2e %.0f"%d['e2e']['value'], "frac %.3f"%r['frac'], {k:round(x,2) for k,x in r['kernels_ms'].items()})
except Exception as e:
    print("$v k=$k FAILED", e); print(open("gpurun_out/bench_var_${v}_k$k.err").read()[-800:])
PY
done
done
