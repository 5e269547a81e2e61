#!/bin/bash
# One GPU-box visit: parity tests, the bench line, the ncu launch list of the same bench command and
# one full capture of the fused kernel.  Outputs under gpurun_out/.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 50 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
tail -c 600 gpurun_out/bench.json
B="python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu --no-alt"
timeout 600 $B > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/ncu_launch.log 2>&1
echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:stream_kernel -s 5 -c 1 -f -o gpurun_out/prof_bench_f64 $B > gpurun_out/ncu_bench_full.log 2>&1
echo "ncu bench full rc=$?"
for dt in f64 f32; do
P="python tools/first_light.py --prof $dt"
timeout 600 $P > gpurun_out/plain_prof_$dt.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:stream_kernel -s 4 -c 1 -f -o gpurun_out/prof_$dt $P > gpurun_out/ncu_prof_$dt.log 2>&1
echo "ncu full $dt rc=$?"
done
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 900 python tools/bench_configs.py > gpurun_out/configs.log 2>&1; echo "configs rc=$?"; grep -v "^{" gpurun_out/configs.log
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; tail -c 400 gpurun_out/bench_ref.json
