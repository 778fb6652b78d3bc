#!/bin/bash
# r03u: idle lanes of the branch-free bucket step aim their atomics at ONE word instead of one word per lane
for v in shipped onespare; do
  if [ $v = shipped ]; then unset MTD_LIB; else export MTD_LIB=$PWD/mazetd_b200/variants/libmazetd_b200_$v.so; fi
  timeout -k 10 200 python bench.py --no-configs --no-e2e --no-cpu --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$v ms/step', round(j['ms_per_step'],3), 'wp_ms', round(j['roofline']['avg_launch_ms'],3))"
  for c in C5 C4; do timeout -k 10 200 python tools/async_time.py $c 65536 2>&1 | grep relax_ms | cut -c1-200; done
done
