#!/bin/bash
# parity suite + bench of the kernel variants.  Usage: gpurun --timeout 900 -- 'bash tools/gpu_variants.sh <tag>'
tag=${1:-v}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$tag.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_$tag.log
for gw in 1 2 4; do for wk in 1 2; do
  echo "gw=$gw walker=$wk"
  timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e --gw $gw --walker $wk 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: j = json.loads(l)
    except Exception: print(l.rstrip()); continue
    print('  ms/step', round(j['ms_per_step'],3), 'wp_ms', round(j['roofline']['avg_launch_ms'],3), 'mapgen_ms', round(j['roofline']['other_kernels']['k_mapgen_grid_avg_ms'],3))
"
done; done | tee gpurun_out/variants_$tag.log
