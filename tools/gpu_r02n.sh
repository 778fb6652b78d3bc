#!/bin/bash
# 2-GPU check: the driver's launch line for bench.py, the peer-coupled stripe test, plain D2H at 2 ranks
mkdir -p gpurun_out
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
timeout -k 10 900 $RUN bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_r02n_n2.json 2> gpurun_out/bench_r02n_n2.err; echo "bench n2 rc=$?"; tail -c 1200 gpurun_out/bench_r02n_n2.err
timeout -k 10 300 $RUN bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > gpurun_out/bench_ref_r02n_n2.json 2>> gpurun_out/bench_r02n_n2.err; echo "ref n2 rc=$?"
timeout -k 10 600 python -m pytest tests -m gpu -x -q -k "peer or stripes" > gpurun_out/pytest_r02n.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_r02n.log | cut -c1-250
timeout -k 10 120 $RUN tools/d2h_bench.py > gpurun_out/d2h_n2_r02n.json 2>&1; tail -1 gpurun_out/d2h_n2_r02n.json
python - <<'PY'
import json
try:
    j = json.loads(open("gpurun_out/bench_r02n_n2.json").read().strip().splitlines()[-1])
    print("C2 weak ms", j["ms_per_step"], "value", j["value"], "strong", j.get("strong"))
    print("e2e", j.get("e2e"))
    for k, v in j.get("configs", {}).items():
        if isinstance(v, dict):
            print(k, {a: v.get(a) for a in ("ms", "value", "parity", "peer_parity", "error", "global_rounds")}, (v.get("roofline") or {}).get("frac"))
            if k == "C3": print("   ", {m: {a: x.get(a) for a in ("wall_s", "value", "collectives_per_rank")} for m, x in v.get("modes", {}).items()})
except Exception as e:
    print("parse failed", e)
PY
