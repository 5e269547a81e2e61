#!/bin/bash
# Multi-GPU visit (gpurun --gpus N): the GPU test suite (sharded NCCL / peer-window tests, devices= keyword), the bench
# under torchrun, host->device bandwidth per rank.
N=${1:-2}
set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus_n$N.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_n$N.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_n$N.log
tail -5 gpurun_out/pytest_gpu_n$N.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"
tail -c 2500 gpurun_out/bench_n$N.json
for n in 1 2 4 8; do
  [ $n -le $N ] || continue
  if [ $n = 1 ]; then timeout 300 python tools/h2d_bw.py >> gpurun_out/h2d_bw_n$N.txt 2>&1
  else timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29612 tools/h2d_bw.py 2>/dev/null | grep '^{' >> gpurun_out/h2d_bw_n$N.txt; fi
done
cat gpurun_out/h2d_bw_n$N.txt
