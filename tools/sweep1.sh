#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 \
  bench.py --gpus 8 --steps 100 --warmup 5 --no-cpu > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err
echo "bench n8 rc=$?"
CNMF_PEER=0 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 \
  bench.py --gpus 8 --steps 100 --warmup 5 --no-cpu --no-e2e > gpurun_out/bench_n8_nccl.json 2> gpurun_out/bench_n8_nccl.err
echo "bench n8 nccl rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29514 \
  bench.py --gpus 4 --steps 100 --warmup 5 --no-cpu --no-e2e > gpurun_out/bench_n4.json 2> gpurun_out/bench_n4.err
echo "bench n4 rc=$?"
