#!/bin/bash
# multi-GPU check (N = number of GPUs of the box): the driver's launch line for bench.py, plain D2H with and without NUMA binding
N=${1:-2}
mkdir -p gpurun_out
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
nvidia-smi topo -m > gpurun_out/r02y_topo_n$N.txt 2>&1; lscpu | grep -E "NUMA|Socket|^CPU\(s\)" >> gpurun_out/r02y_topo_n$N.txt
timeout -k 10 120 $RUN tools/d2h_bench.py > gpurun_out/r02y_d2h_n${N}_bound.json 2>&1; tail -1 gpurun_out/r02y_d2h_n${N}_bound.json
MTD_NO_NUMA_BIND=1 timeout -k 10 120 $RUN tools/d2h_bench.py > gpurun_out/r02y_d2h_n${N}_unbound.json 2>&1; tail -1 gpurun_out/r02y_d2h_n${N}_unbound.json
timeout -k 10 900 $RUN bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r02y_bench_n$N.json 2> gpurun_out/r02y_bench_n$N.err; echo "bench n$N rc=$?"; tail -c 800 gpurun_out/r02y_bench_n$N.err
MTD_NO_NUMA_BIND=1 timeout -k 10 300 $RUN bench.py --gpus $N --steps 10 --warmup 3 --no-configs > gpurun_out/r02y_bench_n${N}_unbound.json 2>> gpurun_out/r02y_bench_n$N.err; echo "bench unbound rc=$?"
python - <<PY
import json
for f in ("gpurun_out/r02y_bench_n$N.json", "gpurun_out/r02y_bench_n${N}_unbound.json"):
    try:
        j = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, "C2 weak ms", j["ms_per_step"], "value", j["value"], "strong", (j.get("strong") or {}).get("ms_per_step"))
        print("  e2e", {k: (j.get("e2e") or {}).get(k) for k in ("value", "ms_per_step", "host_numa")})
        for k, v in (j.get("configs") or {}).items():
            if isinstance(v, dict):
                print("  ", k, {a: v.get(a) for a in ("ms", "value", "parity", "peer_parity", "error")})
    except Exception as e:
        print(f, "parse failed", e)
PY
