#!/bin/bash
# staged device->host results for small maps, asynchronous flag flips: C1 / C3
mkdir -p gpurun_out
timeout -k 10 900 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_r03a.txt 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_r03a.txt | cut -c1-300
timeout -k 10 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_r03a.json 2> gpurun_out/bench_r03a.err; echo "bench rc=$?"; tail -c 400 gpurun_out/bench_r03a.err
python - <<'PY'
import json
j = json.loads(open("gpurun_out/bench_r03a.json").read().strip().splitlines()[-1])
print("C2 ms", j["ms_per_step"], "e2e", j["e2e"]["ms_per_step"])
c = j["configs"]
print("C1", {k: c["C1"].get(k) for k in ("value", "resident_us", "resident_from_scratch_us")}, c["C1"].get("cpu_baseline", {}).get("value"))
print("C3", {k: c["C3"].get(k) for k in ("value", "wall_s", "rebuilds", "trials")})
print("C4", c["C4"].get("ms"), c["C4"].get("parity"), "C5", c["C5"].get("ms"), c["C5"].get("parity"))
PY
