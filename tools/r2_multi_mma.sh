#!/bin/bash
# Multi-GPU visit with the tensor-path kernel as the fp64 default (gpurun --gpus N): the GPU test suite (sharded NCCL /
# peer-window tests, devices= keyword) and the bench under torchrun.
N=${1:-2}
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_mma_n$N.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_mma_n$N.log
tail -5 gpurun_out/pytest_gpu_mma_n$N.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu --no-other > gpurun_out/bench_mma_n$N.json 2> gpurun_out/bench_mma_n$N.err; echo "bench rc=$?"
tail -c 1500 gpurun_out/bench_mma_n$N.json
