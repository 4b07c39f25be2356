cd /root/repo
python -m pytest tests/test_gpu_kernels.py tests/test_gpu_attention_tail.py tests/test_gpu_streaming.py -x -q 2>&1 | tail -4 > gpurun_out/r2_t12.log
python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_size.py -x -q -s 2>&1 | grep "rel-L2\|passed\|failed" >> gpurun_out/r2_t12.log
python tools/gemm_bench.py --m 1920 640 >> gpurun_out/r2_t12.log 2>&1
cat gpurun_out/r2_t12.log
