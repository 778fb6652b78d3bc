#!/bin/bash
# r02q: warps per layer x bucket width on the one-CTA relaxation (variant build with the wide-bucket option), phase split of k_mapgen_grid
mkdir -p gpurun_out
: > gpurun_out/small_gw_delta_r02q.txt
for gw in 2 4; do for d in 0 510 765 1020 1530; do
  echo "== gw $gw delta $d" >> gpurun_out/small_gw_delta_r02q.txt
  MTD_LIB=$PWD/mazetd_b200/variants/libmazetd_b200_wide.so MTD_DEBUG="2=$gw,11=$d" timeout -k 10 120 python bench.py --no-configs --steps 6 --warmup 3 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('  ms/step', round(j['ms_per_step'],3), 'wp_ms', round(j['roofline']['avg_launch_ms'],3))" >> gpurun_out/small_gw_delta_r02q.txt 2>&1
done; done
cat gpurun_out/small_gw_delta_r02q.txt
MTD_LIB=$PWD/mazetd_b200/variants/libmazetd_b200_timing.so timeout -k 10 200 python bench.py --no-configs --steps 3 --warmup 3 > gpurun_out/phases_r02q.txt 2>&1; echo "timing rc=$?"
grep -E "MAPGEN" gpurun_out/phases_r02q.txt | tail -3
