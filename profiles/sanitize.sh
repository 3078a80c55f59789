#!/bin/bash
# compute-sanitizer pass over one small step of every workload (run on the GPU box through gpurun):
#   bash profiles/sanitize.sh <tag>   ->  gpurun_out/<tag>_sanitize_{memcheck,racecheck}.log
TAG=${1:-r2}
mkdir -p gpurun_out
for tool in memcheck racecheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 python profiles/sanitize_driver.py > gpurun_out/${TAG}_sanitize_$tool.log 2>&1
  echo "$tool rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard" gpurun_out/${TAG}_sanitize_$tool.log | tail -5
done
