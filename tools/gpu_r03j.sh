#!/bin/bash
mkdir -p gpurun_out
timeout -k 10 300 python tools/peer_debug.py 2048 2048 16 30 0,256,512,768,1024,1280,1536,1792,2048 > gpurun_out/r03j_a.txt 2>&1; echo "rc=$?"; grep -E "rep|range|tiles|Error|error" gpurun_out/r03j_a.txt | cut -c1-300
timeout -k 10 300 python tools/peer_debug.py 4096 4096 16 30 0,512,1024,1536,2048,2560,3072,3584,4096 > gpurun_out/r03j_b.txt 2>&1; echo "rc=$?"; grep -E "rep|range|tiles|Error|error" gpurun_out/r03j_b.txt | cut -c1-300
timeout -k 10 300 python tools/peer_debug.py 1024 1024 3 30 0,64,128,192,256,320,384,448,512,576,640,704,768,832,896,1024 > gpurun_out/r03j_c.txt 2>&1; echo "rc=$?"; grep -E "rep|range|tiles|Error|error" gpurun_out/r03j_c.txt | cut -c1-300
