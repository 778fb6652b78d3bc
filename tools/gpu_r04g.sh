#!/bin/bash
# r04g: final evidence with the 384-wide buckets: full GPU suite, bench N = 1 + reference arm, ncu launch list
mkdir -p gpurun_out
timeout -k 10 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r04g_pytest_gpu.txt 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r04g_pytest_gpu.txt | cut -c1-300
timeout -k 10 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r04g_bench_n1.json 2> gpurun_out/r04g_bench.err; echo "bench rc=$?"; tail -c 400 gpurun_out/r04g_bench.err
timeout -k 10 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r04g_bench_reference_n1.json 2>> gpurun_out/r04g_bench.err; echo "ref rc=$?"
timeout -k 10 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r04g_ncu_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-configs > gpurun_out/r04g_ncu_bench.log 2>&1; echo "ncu launches rc=$?"
for c in C4 C5; do
timeout -k 10 500 ncu --set full --clock-control none --import-source on -k regex:k_tile_async -s 1 -c 1 -f -o gpurun_out/r04g_async_$c python tools/async_time.py $c 65536 > gpurun_out/r04g_ncu_async_$c.log 2>&1; echo "ncu async $c rc=$?"
done
python - <<'PY'
import json
j = json.loads(open("gpurun_out/r04g_bench_n1.json").read().strip().splitlines()[-1])
print("C2 ms", j["ms_per_step"], "frac", j["roofline"]["frac"], "e2e", j["e2e"]["ms_per_step"])
for k, v in j.get("configs", {}).items():
    if isinstance(v, dict):
        print(k, {a: v.get(a) for a in ("ms", "value", "parity", "error")}, (v.get("roofline") or {}).get("frac"), (v.get("roofline") or {}).get("kernel_avg_ms"))
PY
