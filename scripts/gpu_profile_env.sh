#!/bin/bash
# ncu --set full capture of kernels matching a regex under an environment switch.  usage: gpu_profile_env.sh VAR=VALUE regex outname
mkdir -p gpurun_out
export "$1"
timeout 300 python scripts/profile_pairing.py 16 2 > gpurun_out/plain_prof.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"$2" -s 1 -c 1 -o gpurun_out/$3 -f python scripts/profile_pairing.py 16 2 > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
