#!/bin/bash
# time the tiled large-grid path.  Usage: gpurun --timeout 900 -- 'bash tools/gpu_big.sh <tag> [sizes...]'
tag=${1:-big}; shift
mkdir -p gpurun_out
for s in "${@:-4096}"; do
  timeout 600 python tools/measure_big.py $s 1 >> gpurun_out/big_$tag.log 2>&1
  timeout 600 python tools/measure_big.py $s 16 >> gpurun_out/big_$tag.log 2>&1
done
cat gpurun_out/big_$tag.log
