#!/bin/bash
mkdir -p gpurun_out
for tool in memcheck racecheck synccheck; do
  timeout 600 compute-sanitizer --tool $tool --print-limit 20 python tests/sanitize_mma.py > gpurun_out/sanitize_mma_$tool.log 2>&1; echo "$tool rc=$?"
  tail -6 gpurun_out/sanitize_mma_$tool.log
done
