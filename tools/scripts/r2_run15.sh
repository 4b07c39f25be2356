# in-kernel timelines (debug library): first-operand latency split, per-tile epilogue spans, lean-epilogue chunk stamps and knock-outs
cd /root/repo
export ECHO_B200_LIB=debug
{
python tools/gemm_trace.py --m 1920 --case wo,w2,qkvg,w13
python tools/gemm_trace.py --m 640 --case wo_split,w2_split,qkvg,w13
for d in 1 2 4 7; do echo "### ECHO_GEMM_EPIDBG=$d"; ECHO_GEMM_EPIDBG=$d python tools/gemm_trace.py --m 1920 --case wo; done
} > gpurun_out/r2_trace15.log 2>&1
cat gpurun_out/r2_trace15.log
