#!/bin/bash
# r02x: relaxation of the one-CTA kernel without atomics (checkerboard ownership of the two warps of a layer)
mkdir -p gpurun_out
timeout -k 10 900 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_r02x.txt 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_r02x.txt | cut -c1-400
for gw in 2 1; do
timeout -k 10 300 python bench.py --no-configs --no-e2e --gw $gw --steps 10 --warmup 3 > gpurun_out/bench_r02x_gw$gw.json 2> gpurun_out/bench_r02x.err; echo "bench gw=$gw rc=$?"
python - <<PY
import json
j = json.loads(open("gpurun_out/bench_r02x_gw$gw.json").read().strip().splitlines()[-1])
print("ms/step", j["ms_per_step"], "wp_ms", j["roofline"]["avg_launch_ms"], "frac", j["roofline"]["frac"], j["roofline"]["other_kernels"])
PY
done
