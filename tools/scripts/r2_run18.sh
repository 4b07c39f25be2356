cd /root/repo
for t in 1 2; do echo "== ECHO_GEMM_TAIL=$t"; ECHO_GEMM_TAIL=$t python tools/gemm_bench.py --m 1920 640 2>&1 | grep -v "attn\|norm"; done > gpurun_out/r2_tail18.log 2>&1
ECHO_GEMM_TAIL=2 python -m pytest tests/test_gpu_kernels.py -x -q 2>&1 | tail -2 >> gpurun_out/r2_tail18.log
cat gpurun_out/r2_tail18.log
