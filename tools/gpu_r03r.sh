#!/bin/bash
# r03r: next-hop pass staged through shared memory (k_big_finalize_tiled)
mkdir -p gpurun_out
timeout -k 10 900 python -m pytest tests/test_gpu_parity.py tests/test_adapter_gpu.py -x -q -m gpu -k "tiled or dataflow or c5_full or c4_full or peer or stripes or layers_dev or priority or golden or adapter or directions" > gpurun_out/r03r_pytest.txt 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r03r_pytest.txt | cut -c1-300
timeout -k 10 600 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e > gpurun_out/r03r_bench.json 2> gpurun_out/r03r_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
j = json.loads(open("gpurun_out/r03r_bench.json").read().strip().splitlines()[-1])
for k in ("C4", "C5"):
    r = j["configs"][k]["roofline"]
    print(k, j["configs"][k]["ms"], j["configs"][k]["parity"], "relax", r["kernel_avg_ms"], "next-hop", r["next_hop_pass_avg_ms"], "init", r["init_avg_ms"], "frac", r["frac"])
PY
