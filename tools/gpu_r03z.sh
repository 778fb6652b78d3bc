#!/bin/bash
# r03z: where does a scheduler iteration of the dataflow kernel spend its cycles? (timing build)
mkdir -p gpurun_out
for c in C5 C4; do
  echo "== $c"
  MTD_LIB=$PWD/mazetd_b200/variants/libmazetd_b200_timing.so timeout -k 10 300 python tools/async_time.py $c 65536 2>&1 | grep -E "ASYNCSCHEDPHASE|ASYNCSCHED |relax_ms" | tail -3 | cut -c1-420
done > gpurun_out/r03z_sched_phase.txt 2>&1
cat gpurun_out/r03z_sched_phase.txt
