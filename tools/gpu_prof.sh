#!/bin/bash
# parity suite + one ncu --set full capture of a kernel inside the bench.  Usage: gpurun -- 'bash tools/gpu_prof.sh <tag> <kernel regex> [skip]'
tag=${1:-p}; kern=${2:-k_waypoints_small}; skip=${3:-3}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$tag.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_$tag.log
python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/plain_$tag.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"$kern" -s $skip -c 1 -o gpurun_out/prof_$tag python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_full_$tag.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/plain_$tag.log | cut -c1-600
