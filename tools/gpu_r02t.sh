#!/bin/bash
# r02t: fused generator -- how many warps compute elevations during the relaxation; phase split of the generator stage
mkdir -p gpurun_out
: > gpurun_out/fused_gen_warps_r02t.txt
for gwp in 1 2 3 4 6 8 0; do
  echo "== fused, gen_warps $gwp" >> gpurun_out/fused_gen_warps_r02t.txt
  MTD_DEBUG="12=$gwp" timeout -k 10 120 python bench.py --no-configs --no-e2e --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('  ms/step', round(j['ms_per_step'],3), 'wp_ms', round(j['roofline']['avg_launch_ms'],3))" >> gpurun_out/fused_gen_warps_r02t.txt 2>&1
done
echo "== two kernels" >> gpurun_out/fused_gen_warps_r02t.txt
MTD_DEBUG="11=0" timeout -k 10 120 python bench.py --no-configs --no-e2e --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('  ms/step', round(j['ms_per_step'],3), 'wp_ms', round(j['roofline']['avg_launch_ms'],3), j['roofline']['other_kernels'])" >> gpurun_out/fused_gen_warps_r02t.txt 2>&1
cat gpurun_out/fused_gen_warps_r02t.txt
for gwp in 4 2; do
MTD_DEBUG="12=$gwp" MTD_LIB=$PWD/mazetd_b200/variants/libmazetd_b200_timing.so timeout -k 10 200 python bench.py --no-configs --no-e2e --steps 3 --warmup 3 > gpurun_out/phases_r02t_$gwp.txt 2>&1; echo "timing rc=$?"
grep -E "PHASES" gpurun_out/phases_r02t_$gwp.txt | head -2
done
timeout -k 10 300 python -m pytest tests -x -q -m gpu -k "mapgen" > gpurun_out/pytest_r02t.txt 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_r02t.txt
