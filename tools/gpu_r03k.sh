#!/bin/bash
mkdir -p gpurun_out
timeout -k 10 300 python tools/peer_debug.py 4096 4096 16 30 0,512,1024,1536,2048,2560,3072,3584,4096 > gpurun_out/r03k_b.txt 2>&1; echo "rc=$?"; grep -E "rep|range|tiles|Error|error" gpurun_out/r03k_b.txt | cut -c1-200
timeout -k 10 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "peer or stripes" > gpurun_out/r03k_pytest_peer.txt 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r03k_pytest_peer.txt | cut -c1-300
