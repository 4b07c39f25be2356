#!/bin/bash
cd /root/repo
for sp in 0 2 3; do
  echo "== ECHO_ATTN_SPLITS=$sp" >> gpurun_out/r2_t52.log
  ECHO_ATTN_SPLITS=$sp timeout 200 python tools/gemm_bench.py --m 640 --only attn,seq >> gpurun_out/r2_t52.log 2>&1
done
