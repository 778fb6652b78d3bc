#!/bin/bash
# r03t: warps per layer in the one-CTA relaxation after the branch-free step (idle lanes now cost atomics)
for gw in 2 1 4; do
timeout -k 10 200 python bench.py --no-configs --no-e2e --no-cpu --gw $gw --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('gw $gw ms/step', round(j['ms_per_step'],3), 'wp_ms', round(j['roofline']['avg_launch_ms'],3), j['roofline']['other_kernels'])"
done
