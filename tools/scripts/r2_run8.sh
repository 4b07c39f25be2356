# main-loop experiments: second TMA producer warp for the B operand, ring depth
cd /root/repo
for cfg in "0 10" "1 10" "0 4" "0 3" "1 4"; do
  set -- $cfg
  echo "== ECHO_GEMM_SPLITB=$1 ECHO_GEMM_STAGES=$2" >> gpurun_out/r2_t8_mainloop.log
  ECHO_GEMM_SPLITB=$1 ECHO_GEMM_STAGES=$2 python tools/gemm_bench.py --m 1920 640 --only qkvg,wo,w13,w2,wo/s3,w2/s3,seq >> gpurun_out/r2_t8_mainloop.log 2>&1
done
ECHO_GEMM_SPLITB=1 python -m pytest tests/test_gpu_kernels.py -x -q 2>&1 | tail -3 >> gpurun_out/r2_t8_mainloop.log
cat gpurun_out/r2_t8_mainloop.log
