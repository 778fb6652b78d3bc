#!/bin/bash
# r04b: final evidence for the shipped library: full GPU suite, bench with all legs + reference arm, ncu launch list of the headline step,
# command, full captures of k_waypoints_small (C2) and k_tile_async (C4, C5)
mkdir -p gpurun_out
timeout -k 10 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r04b_pytest_gpu.txt 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r04b_pytest_gpu.txt | cut -c1-300
timeout -k 10 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r04b_bench_n1.json 2> gpurun_out/r04b_bench.err; echo "bench rc=$?"; tail -c 600 gpurun_out/r04b_bench.err
timeout -k 10 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r04b_bench_reference_n1.json 2>> gpurun_out/r04b_bench.err; echo "ref rc=$?"
timeout -k 10 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r04b_ncu_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-configs > gpurun_out/r04b_ncu_bench.log 2>&1; echo "ncu launches rc=$?"
timeout -k 10 400 ncu --set full --clock-control none --import-source on -k regex:k_waypoints_small -s 3 -c 1 -f -o gpurun_out/r04b_wp python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-configs > gpurun_out/r04b_ncu_wp.log 2>&1; echo "ncu wp rc=$?"
timeout -k 10 400 ncu --set full --clock-control none --import-source on -k regex:k_mapgen_grid -s 3 -c 1 -f -o gpurun_out/r04b_mapgen python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-configs > gpurun_out/r04b_ncu_mg.log 2>&1; echo "ncu mapgen rc=$?"
for c in C4 C5; do
MTD_NB=-1 MTD_DEPTH=2 timeout -k 10 500 ncu --set full --clock-control none --import-source on -k regex:k_tile_async -s 1 -c 1 -f -o gpurun_out/r04b_async_$c python tools/async_time.py $c 65536 > gpurun_out/r04b_ncu_async_$c.log 2>&1; echo "ncu async $c rc=$?"; tail -2 gpurun_out/r04b_ncu_async_$c.log | cut -c1-300
done
ls -la gpurun_out/*.ncu-rep
python - <<'PY'
import json
try:
    j = json.loads(open("gpurun_out/r04b_bench_n1.json").read().strip().splitlines()[-1])
    print("C2 ms", j["ms_per_step"], "frac", j["roofline"]["frac"], "e2e", j["e2e"]["ms_per_step"] if j.get("e2e") else None)
    for k, v in j.get("configs", {}).items():
        if isinstance(v, dict):
            print(k, {a: v.get(a) for a in ("ms", "value", "parity", "error")}, (v.get("roofline") or {}).get("frac"), (v.get("e2e") or {}).get("ms"), (v.get("cpu_baseline") or {}).get("value"))
except Exception as e:
    print("parse failed", e)
PY
