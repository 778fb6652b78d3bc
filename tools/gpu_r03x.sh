#!/bin/bash
# r03x: revisits continue on the frame of the worker that visited the tile last (dataflow kernel): parity, then A/B
mkdir -p gpurun_out
timeout -k 10 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "tiled or dataflow or c5_full or c4_full or peer or stripes or layers_dev or priority or golden" > gpurun_out/r03x_pytest.txt 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r03x_pytest.txt | cut -c1-300
: > gpurun_out/r03x_cont_ab.txt
for c in C5 C4; do for k in 0 1; do
  MTD_CONT=$k timeout -k 10 200 python tools/async_time.py $c 65536 2>&1 | grep relax_ms >> gpurun_out/r03x_cont_ab.txt
done; done
cut -c1-220 gpurun_out/r03x_cont_ab.txt
