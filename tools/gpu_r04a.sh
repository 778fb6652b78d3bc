#!/bin/bash
# r04a: every runnable pending entry handed out in one step when the free mailboxes suffice (dataflow scheduler)
mkdir -p gpurun_out
timeout -k 10 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "tiled or dataflow or c5_full or c4_full or peer or stripes or layers_dev or priority or golden" > gpurun_out/r04a_pytest.txt 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r04a_pytest.txt | cut -c1-300
for c in C5 C4; do
  timeout -k 10 200 python tools/async_time.py $c 65536 2>&1 | grep relax_ms | cut -c1-200
  MTD_LIB=$PWD/mazetd_b200/variants/libmazetd_b200_timing.so timeout -k 10 300 python tools/async_time.py $c 65536 2>&1 | grep -E "ASYNCSCHEDPHASE|ASYNCCHAIN" | tail -2 | cut -c1-330
done > gpurun_out/r04a_fast_handout.txt 2>&1
cat gpurun_out/r04a_fast_handout.txt
