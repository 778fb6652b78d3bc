#!/bin/bash
# r04j (8 GPUs): bench at N = 8 with the final library, as the driver launches it
mkdir -p gpurun_out
N=8
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
timeout -k 10 600 $RUN bench.py --gpus $N --steps 10 --warmup 3 --no-cpu > gpurun_out/r04j_bench_n$N.json 2> gpurun_out/r04j_bench_n$N.err; echo "bench n$N rc=$?"; tail -c 300 gpurun_out/r04j_bench_n$N.err
python - <<'PY'
import json
j = json.loads(open("gpurun_out/r04j_bench_n8.json").read().strip().splitlines()[-1])
print("C2 weak ms", j["ms_per_step"], "value", j["value"], "strong", (j.get("strong") or {}).get("ms_per_step"), "e2e", (j.get("e2e") or {}).get("ms_per_step"))
for k, v in (j.get("configs") or {}).items():
    if isinstance(v, dict):
        print("  ", k, {a: v.get(a) for a in ("ms", "ms_all", "value", "parity", "parity_detail", "peer_parity", "error")}, (v.get("roofline") or {}).get("kernel"))
PY
