#!/bin/bash
mkdir -p gpurun_out
for d in 0 510 765 1020; do
  echo "== small delta $d"
  MTD_SMALL_DELTA=$d timeout -k 10 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --no-configs 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: j = json.loads(l)
    except Exception: continue
    print('  ms/step', round(j['ms_per_step'],3), 'wp_ms', round(j['roofline']['avg_launch_ms'],3), 'frac', round(j['roofline']['frac'],4))
"
done > gpurun_out/small_delta_r02o.log 2>&1
cat gpurun_out/small_delta_r02o.log
MTD_SMALL_DELTA=765 timeout -k 10 600 python -m pytest tests -m gpu -x -q -k "waypoints or mapgen or smoke or adapter or incremental or scripted or golden or directions" > gpurun_out/pytest_r02o.log 2>&1; echo "pytest(765) rc=$?"; tail -4 gpurun_out/pytest_r02o.log | cut -c1-250
