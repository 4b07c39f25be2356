# where the first sections epilogue of a launch spends its time; per-kernel times at the streaming-block sizes
cd /root/repo
{
ECHO_B200_LIB=debug python tools/gemm_trace.py --m 640 --case qkvg
ECHO_B200_LIB=debug python tools/gemm_trace.py --m 1920 --case qkvg
python tools/gemm_bench.py --m 160 480 --seq 160
} > gpurun_out/r2_trace21.log 2>&1
cat gpurun_out/r2_trace21.log
