#!/bin/bash
# Round visit: parity suite, bench (both arms), ncu launch list + full capture of the bench kernels, the other
# configurations on one GPU, and a full capture of the large-grid kernel.  Usage: gpurun --timeout 1800 -- 'bash tools/gpu_round2.sh <tag>'
tag=${1:-r01}
bash tools/gpu_round.sh $tag
timeout 600 python tools/bench_configs.py C5 4096 3 > gpurun_out/configs_$tag.jsonl 2> gpurun_out/configs_$tag.err
timeout 600 python tools/bench_configs.py C4 16384 3 >> gpurun_out/configs_$tag.jsonl 2>> gpurun_out/configs_$tag.err
timeout 300 python tools/bench_game.py 10000 1000000 >> gpurun_out/configs_$tag.jsonl 2>> gpurun_out/configs_$tag.err
timeout 300 python bench.py --impl reference --workload C1 --steps 50 >> gpurun_out/configs_$tag.jsonl 2>> gpurun_out/configs_$tag.err
timeout 600 python bench.py --impl reference --workload C5 >> gpurun_out/configs_$tag.jsonl 2>> gpurun_out/configs_$tag.err
cut -c1-330 gpurun_out/configs_$tag.jsonl
timeout 120 python tools/measure_big.py 4096 16 > gpurun_out/plain_big_$tag.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_tile_rounds -c 1 -o gpurun_out/prof_big_$tag python tools/measure_big.py 4096 16 > gpurun_out/ncu_big_$tag.log 2>&1
echo "ncu big rc=$?"; tail -2 gpurun_out/plain_big_$tag.log | cut -c1-250
