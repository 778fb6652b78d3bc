#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "tiled or dataflow or c5_full or c4_full or smoke or golden or towers" > gpurun_out/pytest_r02e.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_r02e.log | cut -c1-300
for c in C5 C4; do
  echo "== timing $c"; MTD_LIB=$PWD/mazetd_b200/variants/libmazetd_b200_timing.so timeout 300 python tools/async_time.py $c 32768,65536,131072 2>&1 | grep -E "ASYNCTIMING|ASYNCSCHED|relax_ms"
  echo "== shipped $c"; timeout 300 python tools/async_time.py $c 16384,32768,65536,131072 2>&1 | grep -E "relax_ms"
done > gpurun_out/variants_r02e.log 2>&1
cut -c1-420 gpurun_out/variants_r02e.log
