#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "tiled or dataflow or c5_full or c4_full or smoke or golden" > gpurun_out/pytest_r02i.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_r02i.log | cut -c1-300
for c in C5 C4; do
  echo "== timing $c"; MTD_NB=-1 MTD_DEPTH=2 MTD_LIB=$PWD/mazetd_b200/variants/libmazetd_b200_timing.so timeout 300 python tools/async_time.py $c 65536 2>&1 | grep -E "ASYNCTIMING|ASYNCSCHED|relax_ms" | tail -4
  echo "== shipped $c"; MTD_NB=-1 MTD_DEPTH=2 timeout 300 python tools/async_time.py $c 32768,65536,131072 2>&1 | grep -E "relax_ms"
done > gpurun_out/variants_r02i.log 2>&1
cut -c1-420 gpurun_out/variants_r02i.log
