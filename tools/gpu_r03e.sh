#!/bin/bash
# r03e: peer-coupled stripes on the dataflow kernel (side by side on one GPU), generator at (512, 2) launch bounds
mkdir -p gpurun_out
timeout -k 10 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "peer_coupled" > gpurun_out/r03e_pytest_peer.txt 2>&1; echo "peer rc=$?"; tail -25 gpurun_out/r03e_pytest_peer.txt | cut -c1-400
timeout -k 10 300 python bench.py --no-configs --no-e2e --steps 10 --warmup 3 > gpurun_out/r03e_bench.json 2> gpurun_out/r03e_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
j = json.loads(open("gpurun_out/r03e_bench.json").read().strip().splitlines()[-1])
print("ms/step", j["ms_per_step"], "wp_ms", j["roofline"]["avg_launch_ms"], j["roofline"]["other_kernels"])
PY
