#!/bin/bash
# r03y: source list requested before the frame (seed phase), shipped + timing build
mkdir -p gpurun_out
for c in C5 C4; do
  MTD_CONT=0 timeout -k 10 200 python tools/async_time.py $c 65536 2>&1 | grep relax_ms | cut -c1-200
  MTD_CONT=0 MTD_LIB=$PWD/mazetd_b200/variants/libmazetd_b200_timing.so timeout -k 10 300 python tools/async_time.py $c 65536 2>&1 | grep -E "ASYNCTIMING" | tail -1 | cut -c1-300
done > gpurun_out/r03y_seed.txt 2>&1
cat gpurun_out/r03y_seed.txt
timeout -k 10 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "tiled or dataflow or c5_full or c4_full or peer or stripes or layers_dev" > gpurun_out/r03y_pytest.txt 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r03y_pytest.txt | cut -c1-200
