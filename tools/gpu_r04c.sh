#!/bin/bash
# r04c: priority window and late looks again, with the parallel hand-out
mkdir -p gpurun_out
for c in C5 C4; do
  timeout -k 10 300 python tools/async_time.py $c 32768,49152,65536,98304,131072 2>&1 | grep relax_ms | cut -c1-200
  MTD_LOOKS=2 timeout -k 10 200 python tools/async_time.py $c 65536 2>&1 | grep relax_ms | cut -c1-200
done > gpurun_out/r04c_window_sweep.txt 2>&1
cat gpurun_out/r04c_window_sweep.txt
