#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "tiled or dataflow or c5_full or smoke" > gpurun_out/pytest_r02g.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_r02g.log | cut -c1-300
for c in C5 C4; do
  for nb in -1 0 8192; do for depth in 1 2; do
    echo "== $c nb=$nb depth=$depth"; MTD_NB=$nb MTD_DEPTH=$depth timeout 300 python tools/async_time.py $c 65536,131072,262144 2>&1 | grep -E "relax_ms"
  done; done
  echo "== timing $c nb=0 depth=1"; MTD_NB=0 MTD_DEPTH=1 MTD_LIB=$PWD/mazetd_b200/variants/libmazetd_b200_timing.so timeout 300 python tools/async_time.py $c 131072 2>&1 | grep -E "ASYNCTIMING|ASYNCSCHED|relax_ms" | tail -4
done > gpurun_out/variants_r02g.log 2>&1
cut -c1-420 gpurun_out/variants_r02g.log
