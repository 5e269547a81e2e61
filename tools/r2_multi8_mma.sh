#!/bin/bash
# 8-GPU visit with the tensor-path kernel as the fp64 default: the bench under torchrun at N = 8 and N = 4 (bench only)
set -x
mkdir -p gpurun_out
for N in 8 4; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2961$N bench.py --gpus $N --steps 20 --warmup 5 --no-cpu --no-other > gpurun_out/bench_mma_n$N.json 2> gpurun_out/bench_mma_n$N.err; echo "bench N=$N rc=$?"
tail -c 600 gpurun_out/bench_mma_n$N.json
done
