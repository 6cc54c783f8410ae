#!/bin/bash
# round-2 GPU visits (one profiling/sanitizer tool per call, B200_PROFILING.md):
#   main      smoke, parity tests, bench (+reference arm), launch list, ncu --set full of the pairing kernels
#   mul       ncu --set full of the scalar-multiplication kernels
#   memcheck / racecheck   compute-sanitizer on the small cases (scripts/sanitizer_cases.py)
mkdir -p gpurun_out
O=gpurun_out
MODE=${1:-main}
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
nproc > $O/nproc.txt
if [ "$MODE" = "main" ]; then
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log
timeout 2400 python -m pytest tests -m gpu -x -q --durations=15 > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
timeout 900 python bench.py --steps 10 --warmup 3 > $O/bench_default.json 2> $O/bench_default.err; echo "rc=$?" >> $O/bench_default.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extra > $O/plain_bench.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extra > $O/ncu_launches.log 2>&1
timeout 300 python scripts/profile_pairing.py 16 2 > $O/plain_prof.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'k_pairs_to_affine|k_g2_lines_sync|k_miller_table_sync|k_final_exp_easy|k_final_exp_hard_sync' -s 5 -c 5 -o $O/prof_pairing -f python scripts/profile_pairing.py 16 2 > $O/ncu_full.log 2>&1
tail -c 300 $O/smoke.log; tail -5 $O/pytest_gpu.log; cut -c1-600 $O/bench_default.json; tail -2 $O/bench_default.err; cut -c1-300 $O/bench_reference.json; tail -3 $O/ncu_full.log
elif [ "$MODE" = "mul" ]; then
timeout 300 python scripts/profile_scalar_mul.py > $O/plain_prof_mul.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'k_jac_mul' -c 8 -o $O/prof_scalar_mul -f python scripts/profile_scalar_mul.py > $O/ncu_full_mul.log 2>&1
tail -3 $O/ncu_full_mul.log
elif [ "$MODE" = "memcheck" ] || [ "$MODE" = "racecheck" ]; then
timeout 300 python scripts/sanitizer_cases.py 8 449 > $O/sanitizer_plain.log 2>&1 &&
timeout 2400 compute-sanitizer --tool $MODE python scripts/sanitizer_cases.py 8 449 > $O/sanitizer_$MODE.log 2>&1; echo "rc=$?" >> $O/sanitizer_$MODE.log
tail -8 $O/sanitizer_$MODE.log
fi
