#!/bin/bash
cd /root/repo
for v in 1 0; do
  echo "== ECHO_GEMM_RESIDL2=$v" >> gpurun_out/r2_t51.log
  ECHO_GEMM_RESIDL2=$v timeout 300 python tools/ae_unit_probe.py 2>&1 | grep "k1 full\|k1 in place\|k1 no snake" >> gpurun_out/r2_t51.log
done
timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_streaming.py -q -k "gemm or ae or decode" 2>&1 | tail -3 >> gpurun_out/r2_t51.log
