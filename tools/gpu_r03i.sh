#!/bin/bash
# r03i (8 GPUs): bench at N = 8 and N = 4 as the driver launches it; per-rank stripe timings at 4 and 8 ranks
mkdir -p gpurun_out
for N in 8 4; do
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
timeout -k 10 600 $RUN bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r03i_bench_n$N.json 2> gpurun_out/r03i_bench_n$N.err; echo "bench n$N rc=$?"; tail -c 300 gpurun_out/r03i_bench_n$N.err
timeout -k 10 300 $RUN tools/peer_time.py 16384 65536 > gpurun_out/r03i_peer_time_n$N.txt 2> gpurun_out/r03i_peer_time_n$N.err; echo "peer_time n$N rc=$?"
grep -E "^\{" gpurun_out/r03i_peer_time_n$N.txt | cut -c1-260
done
python - <<'PY'
import json
for N in (8, 4):
    try:
        j = json.loads(open(f"gpurun_out/r03i_bench_n{N}.json").read().strip().splitlines()[-1])
        print(N, "C2 weak ms", j["ms_per_step"], "value", j["value"], "strong", (j.get("strong") or {}).get("ms_per_step"), "e2e", (j.get("e2e") or {}).get("ms_per_step"))
        for k, v in (j.get("configs") or {}).items():
            if isinstance(v, dict):
                print("  ", k, {a: v.get(a) for a in ("ms", "value", "parity", "peer_parity", "error")}, (v.get("roofline") or {}).get("kernel"))
    except Exception as e:
        print(N, "parse failed", e)
PY
