#!/bin/bash
# r02k: full GPU suite, bench with all legs + reference arm, ncu launch list of the bench command, plain D2H bandwidth
mkdir -p gpurun_out
timeout -k 10 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_r02k.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_r02k.log | cut -c1-300
timeout -k 10 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_r02k.json 2> gpurun_out/bench_r02k.err; echo "bench rc=$?"; tail -c 800 gpurun_out/bench_r02k.err
timeout -k 10 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref_r02k.json 2>> gpurun_out/bench_r02k.err; echo "ref rc=$?"
timeout -k 10 120 python tools/d2h_bench.py > gpurun_out/d2h_n1_r02k.json 2>&1; cat gpurun_out/d2h_n1_r02k.json
timeout -k 10 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02k_ncu_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e > gpurun_out/ncu_bench_r02k.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import json
try:
    j = json.loads(open("gpurun_out/bench_r02k.json").read().strip().splitlines()[-1])
    print("C2 ms", j["ms_per_step"], "frac", j["roofline"]["frac"], "e2e", j["e2e"]["ms_per_step"] if j.get("e2e") else None)
    for k, v in j.get("configs", {}).items():
        if isinstance(v, dict):
            print(k, {a: v.get(a) for a in ("ms", "value", "parity", "error")}, (v.get("roofline") or {}).get("frac"), (v.get("e2e") or {}).get("ms"), (v.get("cpu_baseline") or {}).get("value"))
except Exception as e:
    print("parse failed", e)
PY
