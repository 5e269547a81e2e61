#!/bin/bash
# scratch: one gpurun call's worth of experiments (2 GPUs)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_dist.py -m gpu -x -q > gpurun_out/dist_tests.log 2>&1; echo "dist tests rc=$?" > gpurun_out/s1.log
tail -5 gpurun_out/dist_tests.log >> gpurun_out/s1.log
for peer in 1 0; do
CNMF_PEER=$peer timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus 2 --steps 100 --warmup 5 --no-cpu --no-e2e > gpurun_out/bench_n2_peer$peer.json 2> gpurun_out/bench_n2_peer$peer.err
echo "bench peer=$peer rc=$?" >> gpurun_out/s1.log
done
