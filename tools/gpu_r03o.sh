#!/bin/bash
# r03o: late look at the apron inside a visit (dataflow kernel): parity, then C5 / C4 with 0 / 1 / 2 looks
mkdir -p gpurun_out
timeout -k 10 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "tiled or dataflow or c5_full or c4_full or peer or stripes or layers_dev or priority" > gpurun_out/r03o_pytest.txt 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r03o_pytest.txt | cut -c1-300
: > gpurun_out/r03o_late_looks.txt
for c in C5 C4; do for l in 0 1 2; do
  MTD_LOOKS=$l timeout -k 10 200 python tools/async_time.py $c 65536 2>&1 | grep relax_ms >> gpurun_out/r03o_late_looks.txt
done; done
cat gpurun_out/r03o_late_looks.txt
