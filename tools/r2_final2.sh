#!/bin/bash
# Round 2, one-GPU visit on the code with the fp64 tensor-path kernel as the default: the whole GPU test suite,
# smoke(), the bench line as the driver runs it (--steps 20 --warmup 5) and at the default step count, the CPU
# reference arm, the ncu launch list of the bench command and one full capture of the fused kernel at the bench size.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_s20.json 2> gpurun_out/bench_s20.err; echo "bench s20 rc=$?"
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
tail -c 400 gpurun_out/bench.json
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
B="python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu --no-alt --no-other"
timeout 600 $B > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/ncu_launch.log 2>&1
echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mma_kernel -s 5 -c 1 -f -o gpurun_out/prof_bench_f64_mma $B > gpurun_out/ncu_bench_full.log 2>&1
echo "ncu bench full rc=$?"
CNMF_MMA=0 timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 --no-e2e --no-cpu --no-other > gpurun_out/bench_s20_scalar.json 2> gpurun_out/bench_s20_scalar.err; echo "bench scalar rc=$?"
