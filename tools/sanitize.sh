#!/bin/bash
# compute-sanitizer passes over the warp-specialised kernels (run on the GPU box: gpurun -- bash tools/sanitize.sh).
# Output: gpurun_out/sanitizer_<tool>.log ; summaries are copied into profiles/ by hand.
mkdir -p gpurun_out
for tool in memcheck synccheck racecheck; do
  timeout ${SAN_TIMEOUT:-420} compute-sanitizer --tool $tool --print-limit 20 \
      python tools/sanitize_driver.py > gpurun_out/sanitizer_$tool.log 2>&1
  echo "$tool exit $?" >> gpurun_out/sanitizer_$tool.log
  tail -n 6 gpurun_out/sanitizer_$tool.log
done
