#!/bin/bash
# 8-GPU visit: sharded tests on all 8 ranks, the bench at N = 8 and 4 under torchrun, host->device bandwidth at 1/2/4/8 ranks.
set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus_n8.txt
nvidia-smi topo -m > gpurun_out/topo_n8.txt 2>&1
timeout 1200 python -m pytest tests/test_dist.py tests/test_local_group.py -m gpu -x -q > gpurun_out/pytest_gpu_n8.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_n8.log
tail -5 gpurun_out/pytest_gpu_n8.log
for N in 8 4; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2961$N bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench N=$N rc=$?"
tail -c 600 gpurun_out/bench_n$N.json
done
for n in 1 2 4 8; do
  if [ $n = 1 ]; then timeout 300 python tools/h2d_bw.py >> gpurun_out/h2d_bw_n8.txt 2>&1
  else timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29612 tools/h2d_bw.py 2>/dev/null | grep '^{' >> gpurun_out/h2d_bw_n8.txt; fi
done
cat gpurun_out/h2d_bw_n8.txt
