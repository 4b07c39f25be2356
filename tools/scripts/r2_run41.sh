cd /root/repo
ECHO_GEMM_HYBRID=1 python -m pytest tests/test_gpu_kernels.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -2
for g in 0 1 0 1; do echo "== ECHO_GEMM_HYBRID=$g";
ECHO_GEMM_HYBRID=$g ECHO_B200_LIB=debug python tools/gemm_trace.py --m 1920 --case wo 2>&1 | grep "tile 0"
ECHO_GEMM_HYBRID=$g python tools/gemm_bench.py --m 1920 640 --only wo,w2,wo/s3,w2/s3,seq | grep -v layer
done
