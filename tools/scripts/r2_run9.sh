cd /root/repo
python -m pytest tests/test_gpu_kernels.py tests/test_gpu_streaming.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -4 > gpurun_out/r2_t9.log
python -m pytest tests/test_gpu_parity.py -x -q -k "tiny or full_arch_config1" 2>&1 | tail -3 >> gpurun_out/r2_t9.log
python tools/gemm_bench.py --m 1920 640 >> gpurun_out/r2_t9.log 2>&1
ECHO_B200_LIB=debug python tools/gemm_trace.py --m 640 --case wo_split,w2_split >> gpurun_out/r2_t9.log 2>&1
ECHO_B200_LIB=debug python tools/gemm_trace.py --m 1920 --case wo,w2 >> gpurun_out/r2_t9.log 2>&1
cat gpurun_out/r2_t9.log
