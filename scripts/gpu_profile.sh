#!/bin/bash
# one ncu --set full capture of the pairing kernels (after the plain run exits 0)
mkdir -p gpurun_out
timeout 300 python scripts/profile_pairing.py 16 2 > gpurun_out/plain_prof.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'k_pairs_to_affine|k_g2_lines_sync|k_miller_table_sync|k_final_exp_easy|k_final_exp_hard_sync' -s 5 -c 5 -o gpurun_out/prof_pairing -f python scripts/profile_pairing.py 16 2 > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
