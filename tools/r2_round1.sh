#!/bin/bash
# Round 2, GPU visit 1: parity suite of the restored + hardened tree, the three microbenchmarks, the bench line
# (with the whole chi2 trace for profiles/chi2_trace_n1.json), a few tuning-build sweeps and one full ncu
# capture (with source) of the fused kernel at 1 M rows.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
for b in r2_lds r2_const r2_tcgen05; do
  timeout 120 tools/micro/_bin/$b > gpurun_out/micro_$b.txt 2>&1; echo "$b rc=$?"
done
tail -3 gpurun_out/micro_r2_tcgen05.txt
timeout 1200 python bench.py --steps 150 --warmup 5 --dump-trace > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
tail -c 1500 gpurun_out/bench.json
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; tail -c 600 gpurun_out/bench_ref.json
CNMF_LIB=cnmf_b200/libcnmf_b200_tune.so timeout 600 python tools/first_light.py --sweep "f64:;f64:G=16,RP1=2;f64:NT=4,NOPROD=2;f64:ST=4;f64:ST=3;f32:;f32:G=16" > gpurun_out/sweep.txt 2>&1; echo "sweep rc=$?"
cat gpurun_out/sweep.txt
P="python tools/first_light.py --prof f64"
timeout 600 $P > gpurun_out/plain_prof_f64.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:stream_kernel -s 4 -c 1 -f -o gpurun_out/prof_f64 $P > gpurun_out/ncu_prof_f64.log 2>&1
echo "ncu full f64 rc=$?"
ls -la gpurun_out
