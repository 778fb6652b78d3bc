#!/bin/bash
# quick GPU visit: parity suite (bounded) + short bench.  Usage: gpurun --timeout 900 -- 'bash tools/gpu_quick.sh <tag> [bench args]'
tag=${1:-q}; shift
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$tag.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_$tag.log
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu "$@" > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err; echo "bench rc=$?"
cat gpurun_out/bench_$tag.json; tail -5 gpurun_out/bench_$tag.err
