#!/bin/bash
# r03p: A/B of the dataflow kernel before / after the late look (same box, same script)
mkdir -p gpurun_out
: > gpurun_out/r03p_ab.txt
for c in C5 C4; do
  echo "== $c previous build" >> gpurun_out/r03p_ab.txt
  MTD_LIB=$PWD/mazetd_b200/variants/libmazetd_b200_prev.so timeout -k 10 200 python tools/async_time.py $c 65536 2>&1 | grep relax_ms >> gpurun_out/r03p_ab.txt
  for l in 0 1; do
  echo "== $c current build, looks $l" >> gpurun_out/r03p_ab.txt
  MTD_LOOKS=$l timeout -k 10 200 python tools/async_time.py $c 65536 2>&1 | grep relax_ms >> gpurun_out/r03p_ab.txt
  done
done
cut -c1-200 gpurun_out/r03p_ab.txt
nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,power.draw,temperature.gpu --format=csv
