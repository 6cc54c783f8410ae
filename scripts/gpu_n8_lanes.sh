#!/bin/bash
# N = 8 end-to-end A/B over the pipeline lane count of bn254_pairing_many (host buffers).
mkdir -p gpurun_out
for l in 4 8 6 4 8; do
  BN254_PIPELINE=$l timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus 8 --steps 5 --warmup 3 --no-extra --no-cpu-baseline 2> gpurun_out/n8_lanes$l.err | grep '^{' | tail -1 > gpurun_out/n8_lanes$l.json
  python - <<PY
import json
d = json.load(open("gpurun_out/n8_lanes$l.json"))
print("lanes $l: value %.3f M/s  e2e %.3f M/s" % (d["value"] / 1e6, d["e2e"]["value"] / 1e6))
PY
done
