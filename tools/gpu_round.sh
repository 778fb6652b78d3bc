#!/bin/bash
# One GPU-box visit: parity suite, bench (both arms), ncu launch list and one full capture of the top kernel.
# Usage (here): gpurun --timeout 1500 -- 'bash tools/gpu_round.sh <tag>'
tag=${1:-r01}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv -lms 500 > gpurun_out/clocks_$tag.csv &
SMI=$!
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$tag.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/pytest_$tag.log
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_$tag.json 2> gpurun_out/bench_ref_$tag.err; echo "ref rc=$?"
python bench.py --steps 10 --warmup 3 > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err; echo "bench rc=$?"
cat gpurun_out/bench_$tag.json
kill $SMI
python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/plain_$tag.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_$tag.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_launches_$tag.log 2>&1
echo "ncu launches rc=$?"
python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/plain2_$tag.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_waypoints_small|k_mapgen_grid' -s 6 -c 2 -o gpurun_out/prof_$tag python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_full_$tag.log 2>&1
echo "ncu full rc=$?"
ls -la gpurun_out
