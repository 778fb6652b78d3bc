#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "towers" > gpurun_out/pytest_r02d.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_r02d.log | cut -c1-300
for c in C5 C4; do
  echo "== timing $c"; MTD_LIB=$PWD/mazetd_b200/variants/libmazetd_b200_timing.so timeout 300 python tools/async_time.py $c 65536 2>&1 | grep -E "ASYNCTIMING|ASYNCSCHED|relax_ms" | tail -4
done > gpurun_out/variants_r02d.log 2>&1
cut -c1-420 gpurun_out/variants_r02d.log
