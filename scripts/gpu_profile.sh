#!/bin/bash
# one ncu --set full capture of the two pairing kernels (after the plain run exits 0)
mkdir -p gpurun_out
timeout 300 python scripts/profile_pairing.py 16 2 > gpurun_out/plain_prof.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'k_miller_many|k_final_exp_many' -s 2 -c 2 -o gpurun_out/prof_pairing -f python scripts/profile_pairing.py 16 2 > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
