#!/bin/bash
# r04f: bucket width of the tile engine again, now that the chain of visits (not throughput) bounds the update
mkdir -p gpurun_out
for c in C5 C4; do for v in bd255 bd384 bd512 bd768; do
  if [ $v = shipped ]; then unset MTD_LIB; else export MTD_LIB=$PWD/mazetd_b200/variants/libmazetd_b200_$v.so; fi
  echo -n "$v " ; timeout -k 10 200 python tools/async_time.py $c 65536 2>&1 | grep relax_ms | cut -c1-200
done; done > gpurun_out/r04f_bucket_width.txt 2>&1
cat gpurun_out/r04f_bucket_width.txt
