# sections epilogue: hoisted address arithmetic + sum of squares shared between partner warps
cd /root/repo
python -m pytest tests/test_gpu_kernels.py tests/test_gpu_parity.py -x -q 2>&1 | tail -3
{
ECHO_B200_LIB=debug python tools/gemm_trace.py --m 640 --case qkvg
ECHO_B200_LIB=debug python tools/gemm_trace.py --m 1920 --case qkvg
python tools/gemm_bench.py --m 1920 640 --only qkvg,seq
} > gpurun_out/r2_trace23.log 2>&1
cat gpurun_out/r2_trace23.log
