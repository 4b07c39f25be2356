#!/bin/bash
# stream-K (impl 3) at the 640-token shapes: does balancing the ragged round beat the narrow tail there?
cd /root/repo
echo "== ECHO_STREAM_K=1 ECHO_GEMM_TAIL=0" >> gpurun_out/r2_t48.log
ECHO_GEMM_TAIL=0 ECHO_STREAM_K=1 timeout 200 python tools/gemm_bench.py --m 640 1920 --only qkvg >> gpurun_out/r2_t48.log 2>&1
echo "== ECHO_STREAM_K=0 ECHO_GEMM_TAIL=0" >> gpurun_out/r2_t48.log
ECHO_GEMM_TAIL=0 ECHO_STREAM_K=0 timeout 200 python tools/gemm_bench.py --m 640 1920 --only qkvg >> gpurun_out/r2_t48.log 2>&1
