#!/bin/bash
# r03w: how long is the chain of dependent visits? (timing build)
mkdir -p gpurun_out
for c in C5 C4; do for k in 0 1; do
  echo "== $c cont=$k"
  MTD_CONT=$k MTD_LIB=$PWD/mazetd_b200/variants/libmazetd_b200_timing.so timeout -k 10 300 python tools/async_time.py $c 65536 2>&1 | grep -E "ASYNCCHAIN|ASYNCTIMING|relax_ms" | tail -3 | cut -c1-420
done; done > gpurun_out/r03w_chain.txt 2>&1
cat gpurun_out/r03w_chain.txt
