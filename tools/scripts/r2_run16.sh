# scheduler divisions hoisted above the dependency wait: kernel tests, timelines, per-projection timings
cd /root/repo
python -m pytest tests/test_gpu_kernels.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -3
{
ECHO_B200_LIB=debug python tools/gemm_trace.py --m 1920 --case wo,qkvg
ECHO_B200_LIB=debug python tools/gemm_trace.py --m 640 --case wo_split,w13
python tools/gemm_bench.py --m 1920 640
} > gpurun_out/r2_trace16.log 2>&1
cat gpurun_out/r2_trace16.log
