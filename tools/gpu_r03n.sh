#!/bin/bash
# r03n (8-GPU node): plain pinned device->host copies at 1 / 2 / 4 / 8 ranks: the node's host ceiling next to the e2e leg of the bench
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r03n_topo_n8.txt 2>&1; lscpu | grep -E "NUMA|Socket|^CPU\(s\)|Model name" >> gpurun_out/r03n_topo_n8.txt
: > gpurun_out/r03n_d2h_node8.jsonl
for N in 1 2 4 8; do
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29535"
timeout -k 10 120 $RUN tools/d2h_bench.py 2>/dev/null | grep -E "^\{" >> gpurun_out/r03n_d2h_node8.jsonl
MTD_NO_NUMA_BIND=1 timeout -k 10 120 $RUN tools/d2h_bench.py 2>/dev/null | grep -E "^\{" | sed 's/^{/{"unbound": true, /' >> gpurun_out/r03n_d2h_node8.jsonl
done
cat gpurun_out/r03n_d2h_node8.jsonl | cut -c1-400; head -12 gpurun_out/r03n_topo_n8.txt
