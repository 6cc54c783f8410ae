#!/bin/bash
# multi-GPU visit (gpurun --gpus N): parity tests with real peer copies, and the NCCL-sharded product
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_multi.log 2>&1; echo "pytest rc=$? on $N GPUs" >> gpurun_out/pytest_gpu_multi.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29571 scripts/dist_check.py 4096 > gpurun_out/dist_check.log 2>&1; echo "dist_check rc=$?" >> gpurun_out/dist_check.log
tail -3 gpurun_out/pytest_gpu_multi.log; grep -E "ok:|rc=|Error|error" gpurun_out/dist_check.log | tail -12
