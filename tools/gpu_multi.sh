#!/bin/bash
# Multi-GPU visit: bench.py (C2, maps sharded per GPU) and tools/bench_configs.py (C4 row stripes, C5 layers) at N ranks.
# Usage (here): gpurun --gpus N --timeout 1500 -- 'bash tools/gpu_multi.sh <tag> N [c4size]'
tag=${1:-multi}; N=${2:-2}; C4=${3:-16384}
mkdir -p gpurun_out
out=gpurun_out/multi_${tag}_n$N.jsonl
: > $out
if [ "$N" = 1 ]; then RUN="python"; else RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"; fi
timeout 600 $RUN bench.py --gpus $N --steps 10 --warmup 3 --no-cpu >> $out 2> gpurun_out/multi_${tag}_n$N.err; echo "bench rc=$?"
timeout 900 $RUN tools/bench_configs.py C5 4096 3 >> $out 2>> gpurun_out/multi_${tag}_n$N.err; echo "C5 rc=$?"
timeout 1200 $RUN tools/bench_configs.py C4 $C4 3 >> $out 2>> gpurun_out/multi_${tag}_n$N.err; echo "C4 rc=$?"
cut -c1-420 $out
tail -5 gpurun_out/multi_${tag}_n$N.err
