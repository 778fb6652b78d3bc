#!/bin/bash
# r02s: generator fused into k_waypoints_small (next map's elevations computed during the relaxation), union-find with path halving
mkdir -p gpurun_out
timeout -k 10 900 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_r02s.txt 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_r02s.txt
for f in 1 0; do
MTD_DEBUG="11=$f" timeout -k 10 300 python bench.py --no-configs --steps 10 --warmup 3 > gpurun_out/bench_r02s_fused$f.json 2> gpurun_out/bench_r02s.err; echo "bench fused=$f rc=$?"
python - <<PY
import json
j = json.loads(open("gpurun_out/bench_r02s_fused$f.json").read().strip().splitlines()[-1])
print("ms/step", j["ms_per_step"], "wp_ms", j["roofline"]["avg_launch_ms"], "frac", j["roofline"]["frac"], "other", j["roofline"]["other_kernels"], "e2e", j["e2e"]["value"])
PY
done
MTD_LIB=$PWD/mazetd_b200/variants/libmazetd_b200_timing.so timeout -k 10 200 python bench.py --no-configs --steps 3 --warmup 3 > gpurun_out/phases_r02s.txt 2>&1; echo "timing rc=$?"
grep -E "MAPGEN|PHASES" gpurun_out/phases_r02s.txt | head -3
MTD_DEBUG="11=0" MTD_LIB=$PWD/mazetd_b200/variants/libmazetd_b200_timing.so timeout -k 10 200 python bench.py --no-configs --steps 3 --warmup 3 > gpurun_out/phases_r02s_unfused.txt 2>&1; echo "timing rc=$?"
grep -E "MAPGEN|PHASES" gpurun_out/phases_r02s_unfused.txt | head -3
