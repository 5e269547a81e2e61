#!/bin/bash
# A/B of two product builds on one box: libcnmf_b200_old.so (before) vs libcnmf_b200.so (after), plus parity of the new one.
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py -x -q > gpurun_out/pytest_new.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_new.log
tail -4 gpurun_out/pytest_new.log
for rep in 1 2; do
for lib in old new; do
  L=$PWD/cnmf_b200/libcnmf_b200.so; [ $lib = old ] && L=$PWD/cnmf_b200/libcnmf_b200_old.so
  echo "== $lib" >> gpurun_out/ab.txt
  CNMF_LIB=$L timeout 600 python tools/first_light.py --sweep "f64:;f32:" >> gpurun_out/ab.txt 2>&1
  CNMF_LIB=$L timeout 600 python tools/first_light.py --sweep "f32:" --n 1700000 --m 184 --k 8 >> gpurun_out/ab.txt 2>&1
  CNMF_LIB=$L timeout 600 python tools/first_light.py --sweep "f64:" --n 100000 --m 301 --k 4 >> gpurun_out/ab.txt 2>&1
done
done
grep -E "^==|^timing" gpurun_out/ab.txt
