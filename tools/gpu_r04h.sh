#!/bin/bash
# r04h: priority window once more, with the 384-wide buckets
mkdir -p gpurun_out
for c in C5 C4; do
  timeout -k 10 300 python tools/async_time.py $c 32768,49152,65536,98304 2>&1 | grep relax_ms | cut -c1-200
done > gpurun_out/r04h_window.txt 2>&1
cat gpurun_out/r04h_window.txt
