#!/bin/bash
# r03g (2 GPUs): peer-coupled stripes on the dataflow kernel -- parity tests (side by side and two ranks), bench at N = 2
mkdir -p gpurun_out
timeout -k 10 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "peer or stripes" > gpurun_out/r03g_pytest_peer_2gpu.txt 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r03g_pytest_peer_2gpu.txt | cut -c1-300
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
timeout -k 10 900 $RUN bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r03g_bench_n2.json 2> gpurun_out/r03g_bench_n2.err; echo "bench n2 rc=$?"; tail -c 600 gpurun_out/r03g_bench_n2.err
python - <<'PY'
import json
try:
    j = json.loads(open("gpurun_out/r03g_bench_n2.json").read().strip().splitlines()[-1])
    print("C2 weak ms", j["ms_per_step"], "value", j["value"], "strong", (j.get("strong") or {}).get("ms_per_step"))
    for k, v in (j.get("configs") or {}).items():
        if isinstance(v, dict):
            print("  ", k, {a: v.get(a) for a in ("ms", "ms_all", "value", "parity", "peer_parity", "error", "tile_visits_rank0")}, (v.get("roofline") or {}).get("kernel"))
except Exception as e:
    print("parse failed", e)
PY
