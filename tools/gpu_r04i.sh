#!/bin/bash
# r04i: worker warps per SM once more with the final kernel
mkdir -p gpurun_out
for c in C5 C4; do for w in 10 12 14 16; do
  MTD_WORKERS=$w timeout -k 10 200 python tools/async_time.py $c 65536 2>&1 | grep relax_ms | cut -c1-200
done; done > gpurun_out/r04i_workers.txt 2>&1
cat gpurun_out/r04i_workers.txt
