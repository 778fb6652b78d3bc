#!/bin/bash
# r02a: full GPU suite with the shipped library, probe variants of the dataflow kernel, a first bench with all legs
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_r02a.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_r02a.log
for v in timing t512 t2048 t4096; do
  for c in C5 C4; do
    echo "== $v $c"; MTD_LIB=$PWD/mazetd_b200/variants/libmazetd_b200_$v.so timeout 300 python tools/async_time.py $c 65536 2>&1 | grep -E "ASYNCTIMING|relax_ms" | tail -2
  done
done > gpurun_out/variants_r02a.log 2>&1
cat gpurun_out/variants_r02a.log | cut -c1-400
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_r02a.json 2> gpurun_out/bench_r02a.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/bench_r02a.err
python - <<'PY'
import json
try:
    j = json.loads(open("gpurun_out/bench_r02a.json").read().strip().splitlines()[-1])
    print("C2 ms", j["ms_per_step"], "e2e", j["e2e"]["ms_per_step"] if j.get("e2e") else None)
    for k, v in j.get("configs", {}).items():
        print(k, {a: v.get(a) for a in ("ms", "value", "parity", "error")}, (v.get("roofline") or {}).get("frac"), (v.get("e2e") or {}).get("ms"), (v.get("cpu_baseline") or {}).get("value"))
except Exception as e:
    print("parse failed", e)
PY
