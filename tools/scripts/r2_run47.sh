#!/bin/bash
cd /root/repo
for o in 0 1; do
  echo "== ECHO_CHAIN_ORDER=$o" >> gpurun_out/r2_t47.log
  ECHO_CHAIN_ORDER=$o timeout 200 python -m pytest tests/test_gpu_kernels.py -x -q -k "mlp_chain" 2>&1 | tail -2 >> gpurun_out/r2_t47.log
  ECHO_CHAIN_ORDER=$o timeout 300 python tools/gemm_bench.py --m 1920 640 --only mlp2,chain,seq,seqc >> gpurun_out/r2_t47.log 2>&1
  ECHO_CHAIN_ORDER=$o timeout 300 python bench.py --steps 3 --warmup 3 > gpurun_out/r2_t47_bench_$o.json 2> gpurun_out/r2_t47_bench_$o.err
done
ECHO_MLP_CHAIN=0 timeout 300 python bench.py --steps 3 --warmup 3 > gpurun_out/r2_t47_bench_off.json 2> gpurun_out/r2_t47_bench_off.err
