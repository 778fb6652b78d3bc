#!/bin/bash
# r03s: noise without branches around the simplex corners under the (512, 2) launch bounds: parity + headline
mkdir -p gpurun_out
for v in shipped bfnoise; do
  if [ $v = shipped ]; then unset MTD_LIB; else export MTD_LIB=$PWD/mazetd_b200/variants/libmazetd_b200_$v.so; fi
  timeout -k 10 200 python bench.py --no-configs --no-e2e --no-cpu --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$v ms/step', round(j['ms_per_step'],3), 'wp_ms', round(j['roofline']['avg_launch_ms'],3), j['roofline']['other_kernels'])"
done
export MTD_LIB=$PWD/mazetd_b200/variants/libmazetd_b200_bfnoise.so
timeout -k 10 400 python -m pytest tests/test_gpu_parity.py tests/test_adapter_gpu.py -x -q -m gpu -k "mapgen or elevation or sdf or golden or adapter or grid_from" > gpurun_out/r03s_pytest.txt 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r03s_pytest.txt
