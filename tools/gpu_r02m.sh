#!/bin/bash
mkdir -p gpurun_out
timeout -k 10 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_r02m.json 2> gpurun_out/bench_r02m.err; echo "bench rc=$?"; tail -c 600 gpurun_out/bench_r02m.err
python - <<'PY'
import json
j = json.loads(open("gpurun_out/bench_r02m.json").read().strip().splitlines()[-1])
c = j["configs"]
print("C1", {k: c["C1"].get(k) for k in ("value", "resident_us", "resident_from_scratch_us", "incremental_updates", "incremental_trials")}, c["C1"].get("cpu_baseline", {}).get("value"))
print("C3", {k: c["C3"].get(k) for k in ("value", "wall_s", "rebuilds", "trials")})
print("C4", c["C4"].get("ms"), c["C4"].get("parity"), "C5", c["C5"].get("ms"), c["C5"].get("parity"))
PY
