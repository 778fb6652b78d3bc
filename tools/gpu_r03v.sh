#!/bin/bash
# r03v: final evidence of the round for the shipped library: full GPU suite, bench N = 1 with all legs + reference arm, ncu launch list
mkdir -p gpurun_out
timeout -k 10 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r03v_pytest_gpu.txt 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r03v_pytest_gpu.txt | cut -c1-300
timeout -k 10 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r03v_bench_n1.json 2> gpurun_out/r03v_bench.err; echo "bench rc=$?"; tail -c 400 gpurun_out/r03v_bench.err
timeout -k 10 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r03v_bench_reference_n1.json 2>> gpurun_out/r03v_bench.err; echo "ref rc=$?"
timeout -k 10 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r03v_ncu_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-configs > gpurun_out/r03v_ncu_bench.log 2>&1; echo "ncu launches rc=$?"
timeout -k 10 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
python - <<'PY'
import json
j = json.loads(open("gpurun_out/r03v_bench_n1.json").read().strip().splitlines()[-1])
print("C2 ms", j["ms_per_step"], "value", j["value"], "frac", j["roofline"]["frac"], "share", j["roofline"]["share_of_step"], "e2e", j["e2e"]["ms_per_step"], "launches", j["gpu_launches"], "clocks", j["clocks"])
for k, v in j.get("configs", {}).items():
    if isinstance(v, dict):
        print(k, {a: v.get(a) for a in ("ms", "value", "parity", "error")}, (v.get("roofline") or {}).get("frac"), (v.get("roofline") or {}).get("traffic"))
r = json.loads(open("gpurun_out/r03v_bench_reference_n1.json").read().strip().splitlines()[-1])
print("reference", r["value"], r["unit"], r.get("impl"))
PY
