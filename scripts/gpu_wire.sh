#!/bin/bash
# GPU visit for SURVEY 8f N2/N3: parity tests, wire throughput, refreshed config #3/#4 numbers, default bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 900 python scripts/bench_wire.py > gpurun_out/bench_wire.txt 2> gpurun_out/bench_wire.err; echo "rc=$?" >> gpurun_out/bench_wire.err
BN254_SUBGROUP_REF=1 timeout 900 python scripts/bench_wire.py >> gpurun_out/bench_wire.txt 2>> gpurun_out/bench_wire.err; echo "rc=$?" >> gpurun_out/bench_wire.err
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "rc=$?" >> gpurun_out/bench_default.err
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/bench_wire.txt; tail -3 gpurun_out/bench_wire.err; cut -c1-300 gpurun_out/bench_default.json
