#!/bin/bash
cd /root/repo
timeout 600 python -m pytest tests/test_gpu_kernels.py -x -q -k "row_epilogue" 2>&1 | tail -15 > gpurun_out/r2_t56_tests.log
if grep -q "passed" gpurun_out/r2_t56_tests.log && ! grep -q "failed\|error" gpurun_out/r2_t56_tests.log; then
  timeout 300 python tools/ae_unit_probe.py > gpurun_out/r2_t56_probe.log 2>&1
  timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_streaming.py tests/test_gpu_parity.py tests/test_gpu_encode.py -x -q 2>&1 | tail -5 > gpurun_out/r2_t56_suite.log
fi
