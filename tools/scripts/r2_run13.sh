# final-kernel evidence: tests for the advisor fixes, ncu (GEMMs at both step shapes, each HBM-bound kernel), launch list
cd /root/repo
python -m pytest tests/test_gpu_attention_tail.py tests/test_gpu_parity.py -x -q -k "flattening or round_t or race_free or tiny" 2>&1 | tail -3 > gpurun_out/r2_t13.log
python tools/gemm_bench.py --quick --only wo,w2,qkvg,w13,attn --m 1920 > gpurun_out/r2_plain1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"gemm_pair_kernel|attention_tc_kernel" -c 5 -o gpurun_out/r2f_gemm1920 -f python tools/gemm_bench.py --quick --only wo,w2,qkvg,w13,attn --m 1920 > gpurun_out/r2_ncu1.log 2>&1
python tools/gemm_bench.py --quick --only wo/s3,w2/s3,qkvg,w13,attn --m 640 > gpurun_out/r2_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"gemm_pair_kernel|attention_tc_kernel" -c 5 -o gpurun_out/r2f_gemm640 -f python tools/gemm_bench.py --quick --only wo/s3,w2/s3,qkvg,w13,attn --m 640 > gpurun_out/r2_ncu2.log 2>&1
python tools/profile_step.py > gpurun_out/r2_plain3.log 2>&1 && {
for k in rmsnorm_mod_kernel cfg_euler_tail_kernel cast_snake_kernel dwconv_ln_kernel conv_out_tanh_kernel sgemm_kernel; do
  ncu --set full --clock-control none --profile-from-start off -k regex:$k -c 3 -o gpurun_out/r2f_hbm_$k -f python tools/profile_step.py > gpurun_out/r2_ncu3_$k.log 2>&1
done
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 2500 --csv --log-file gpurun_out/r2f_launches_step.csv python tools/profile_step.py > gpurun_out/r2_ncu4.log 2>&1
}
cat gpurun_out/r2_t13.log; ls gpurun_out | grep r2f
