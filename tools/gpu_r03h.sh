#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534"
timeout -k 10 600 $RUN tools/peer_time.py 16384 32768,65536,262144,4294967295 > gpurun_out/r03h_peer_time_n$N.txt 2> gpurun_out/r03h_peer_time_n$N.err; echo "rc=$?"; grep -E "^\{" gpurun_out/r03h_peer_time_n$N.txt | cut -c1-330; tail -3 gpurun_out/r03h_peer_time_n$N.err | cut -c1-300
