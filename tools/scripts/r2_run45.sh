#!/bin/bash
# stream-K as impl 3 / plain schedule as impl 4; conv_out rewrite: order-dependent subset first, then the full suite
cd /root/repo
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_streaming.py -q -k "conv_out or ae or decode or stream" 2>&1 | tail -8 > gpurun_out/r2_t45_subset.log
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r2_t45_full.log
timeout 300 python bench.py --steps 3 --warmup 3 > gpurun_out/r2_t45_bench.json 2> gpurun_out/r2_t45_bench.err
