#!/bin/bash
# parity tests of the tiled path + timing.  Usage: gpurun --timeout 900 -- 'bash tools/gpu_big.sh <tag> [sizes...]'
tag=${1:-big}; shift
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "tiled or golden or more_than_four or fewer_resident or smoke" > gpurun_out/pytest_big_$tag.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_big_$tag.log
rm -f gpurun_out/big_$tag.log
for s in "${@:-4096}"; do
  timeout 600 python tools/measure_big.py $s 1 >> gpurun_out/big_$tag.log 2>&1
  timeout 600 python tools/measure_big.py $s 16 >> gpurun_out/big_$tag.log 2>&1
done
grep -v "^gen" gpurun_out/big_$tag.log | cut -c1-220
