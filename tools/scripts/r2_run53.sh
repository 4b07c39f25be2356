#!/bin/bash
# marginal cost of every launch inside the layer sequence: seq with one launch left out
cd /root/repo
for m in 640 1920; do
  for d in "" norm1 qkvg attn wo norm2 w13 w2; do
    echo "== M=$m drop=$d" >> gpurun_out/r2_t53.log
    timeout 120 python tools/gemm_bench.py --m $m --only seq --drop "$d" 2>&1 | grep seq >> gpurun_out/r2_t53.log
  done
done
