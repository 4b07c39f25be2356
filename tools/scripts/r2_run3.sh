cd /root/repo
for sk in 0 1; do
  echo "== ECHO_STREAM_K=$sk" >> gpurun_out/r2_t3_streamk.log
  ECHO_STREAM_K=$sk python tools/gemm_bench.py --m 640 --only wo,w2,qkvg,w13 >> gpurun_out/r2_t3_streamk.log 2>&1
done
ECHO_STREAM_K=1 python tools/gemm_bench.py --m 1920 --only wo,w2,qkvg,w13 >> gpurun_out/r2_t3_streamk.log 2>&1
# sanitizer passes over the kernel unit tests at small shapes (hand-rolled mbarrier / TMEM / stream-K protocols)
SEL='test_gemm_plain and (128-256-64 or 200-384 or 130-64 or 64-96) or test_gemm_swiglu or test_rmsnorm or test_gemm_sections_norm_rope or test_gemm_generic_epilogue_full'
timeout 700 compute-sanitizer --tool memcheck --log-file gpurun_out/r2_sanitizer_memcheck.log python -m pytest tests/test_gpu_kernels.py -x -q -k "$SEL" > gpurun_out/r2_sanitizer_memcheck_pytest.log 2>&1
echo "memcheck rc=$?" >> gpurun_out/r2_sanitizer_memcheck_pytest.log
timeout 700 compute-sanitizer --tool racecheck --log-file gpurun_out/r2_sanitizer_racecheck.log python -m pytest tests/test_gpu_kernels.py -x -q -k "$SEL" > gpurun_out/r2_sanitizer_racecheck_pytest.log 2>&1
echo "racecheck rc=$?" >> gpurun_out/r2_sanitizer_racecheck_pytest.log
tail -3 gpurun_out/r2_sanitizer_memcheck_pytest.log gpurun_out/r2_sanitizer_racecheck_pytest.log
tail -5 gpurun_out/r2_sanitizer_memcheck.log gpurun_out/r2_sanitizer_racecheck.log
