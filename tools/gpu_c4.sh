#!/bin/bash
# C4 (16384^2 row stripes) at N ranks for several halo-swap budgets. Usage: gpurun --gpus N -- 'bash tools/gpu_c4.sh <tag> N "<budgets>" [size]'
tag=${1:-c4}; N=${2:-2}; budgets=${3:-"0 8"}; size=${4:-16384}
mkdir -p gpurun_out
out=gpurun_out/c4_${tag}_n$N.jsonl
: > $out
for b in $budgets; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 tools/bench_configs.py C4 $size 3 $b >> $out 2> gpurun_out/c4_${tag}_n$N.err; echo "budget $b rc=$?"
done
python - <<PY
import json
for l in open("$out"):
    try: j = json.loads(l)
    except Exception: continue
    print(j["n_gpus"], "gpus budget", j["round_budget"], "rep", j["rep"], "ms %.2f" % j["ms"], "sweeps", j["sweeps"], "launches", j["launches_rank0"], j["dist_checksum"])
PY
tail -3 gpurun_out/c4_${tag}_n$N.err
