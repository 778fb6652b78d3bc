#!/bin/bash
# r02z: dataflow kernel -- how many worker warps per SM are handed items (the bucket step is bound by the SM's shared-memory pipeline)
mkdir -p gpurun_out
: > gpurun_out/r02z_async_workers.txt
for c in C5 C4; do for w in 6 8 10 12 14 16; do
  MTD_WORKERS=$w timeout -k 10 200 python tools/async_time.py $c 65536 2>&1 | grep relax_ms >> gpurun_out/r02z_async_workers.txt
done; done
cat gpurun_out/r02z_async_workers.txt
