#!/bin/bash
# parity tests, then BASELINE configs #3/#4 (scripts/bench_configs.py), optionally under an environment switch for A/B
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 900 python scripts/bench_configs.py > gpurun_out/bench_configs.txt 2>&1; echo "rc=$?" >> gpurun_out/bench_configs.txt
if [ -n "$1" ]; then export "$1"; timeout 900 python scripts/bench_configs.py > gpurun_out/bench_configs_alt.txt 2>&1; fi
cut -c1-260 gpurun_out/bench_configs.txt; if [ -f gpurun_out/bench_configs_alt.txt ]; then cut -c1-260 gpurun_out/bench_configs_alt.txt; fi
