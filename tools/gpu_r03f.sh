#!/bin/bash
mkdir -p gpurun_out
timeout -k 10 200 python tools/peer_debug.py > gpurun_out/r03f_peer_debug.txt 2>&1; echo "rc=$?"; tail -30 gpurun_out/r03f_peer_debug.txt | cut -c1-400
timeout -k 10 200 python tools/peer_debug.py 1024 1024 16 30 0,256,512,768,1024 > gpurun_out/r03f_peer_debug2.txt 2>&1; echo "rc=$?"; tail -30 gpurun_out/r03f_peer_debug2.txt | cut -c1-400
