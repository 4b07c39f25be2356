#!/bin/bash
# chained MLP launch: correctness first (own timeout: a hang must not hold the box), then timings
cd /root/repo
timeout 300 python -m pytest tests/test_gpu_kernels.py -x -q -k "mlp_chain" 2>&1 | tail -15 > gpurun_out/r2_t46_tests.log
rc=$?
if grep -q "passed" gpurun_out/r2_t46_tests.log && ! grep -q "failed" gpurun_out/r2_t46_tests.log; then
  timeout 300 python tools/gemm_bench.py --m 1920 640 --only w13,w2,w2/s3,mlp2,chain,seq,seqc > gpurun_out/r2_t46_bench.log 2>&1
  timeout 200 python tools/gemm_bench.py --m 160 480 --seq 160 --only mlp2,chain,seq,seqc >> gpurun_out/r2_t46_bench.log 2>&1
  timeout 300 python bench.py --steps 3 --warmup 3 > gpurun_out/r2_t46_bench.json 2> gpurun_out/r2_t46_bench.err
  ECHO_MLP_CHAIN=0 timeout 300 python bench.py --steps 3 --warmup 3 > gpurun_out/r2_t46_bench_off.json 2> gpurun_out/r2_t46_bench_off.err
fi
