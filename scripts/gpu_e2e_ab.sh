#!/bin/bash
# A/B of the zero-copy output path: e2e with and without it
mkdir -p gpurun_out
for z in 1 0 1 0; do
  BN254_ZEROCOPY_OUT=$z timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_zc$z.json 2> gpurun_out/bench_zc$z.err
  python - <<PY
import json
d=json.load(open("gpurun_out/bench_zc$z.json")); print("zerocopy=$z value %.0f e2e %.0f" % (d['value'], d['e2e']['value']))
PY
done
