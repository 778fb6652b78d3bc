#!/bin/bash
mkdir -p gpurun_out
for c in C5 C4; do
  echo "== timing $c"; MTD_LIB=$PWD/mazetd_b200/variants/libmazetd_b200_timing.so timeout 300 python tools/async_time.py $c 32768,65536,131072 2>&1 | grep -E "ASYNCTIMING|ASYNCSCHED|relax_ms"
  echo "== shipped $c"; timeout 300 python tools/async_time.py $c 65536 2>&1 | grep -E "relax_ms"
done > gpurun_out/variants_r02c.log 2>&1
cut -c1-420 gpurun_out/variants_r02c.log
