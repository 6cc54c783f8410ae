#!/bin/bash
# first GPU contact: smoke, parity tests, microbench + bench on both library variants, launch list
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
nproc > gpurun_out/nproc.txt
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "rc=$?" >> gpurun_out/bench_default.err
timeout 600 python bench.py --steps 3 --warmup 3 --variant _r128 --no-cpu-baseline > gpurun_out/bench_r128.json 2> gpurun_out/bench_r128.err; echo "rc=$?" >> gpurun_out/bench_r128.err
timeout 600 python bench.py --steps 2 --warmup 3 --batch-log2 18 --no-cpu-baseline > gpurun_out/bench_default_2p18.json 2> gpurun_out/bench_default_2p18.err
timeout 600 python bench.py --steps 2 --warmup 3 --batch-log2 18 --variant _r128 --no-cpu-baseline > gpurun_out/bench_r128_2p18.json 2> gpurun_out/bench_r128_2p18.err
tail -c 600 gpurun_out/smoke.log; tail -5 gpurun_out/pytest_gpu.log; cat gpurun_out/bench_default.json | cut -c1-1500; tail -3 gpurun_out/bench_default.err
