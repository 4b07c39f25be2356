# ncu: (1) full capture of the five GEMM shapes + attention (stall reasons in the epilogues), (2) the HBM-bound kernels of a
# CFG step + plain step + decode at configs[1] shapes (dram bytes / throughput)
cd /root/repo
python tools/gemm_bench.py --quick --only wo,w2,qkvg,w13,attn --m 1920 > gpurun_out/r2_t4_plain1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"gemm_pair_kernel|attention_tc_kernel" -c 5 -o gpurun_out/r2_gemm1920 -f python tools/gemm_bench.py --quick --only wo,w2,qkvg,w13,attn --m 1920 > gpurun_out/r2_t4_ncu1.log 2>&1
python tools/gemm_bench.py --quick --only wo/s3,w2/s3,attn --m 640 > gpurun_out/r2_t4_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"gemm_pair_kernel|attention_tc_kernel" -c 3 -o gpurun_out/r2_gemm640 -f python tools/gemm_bench.py --quick --only wo/s3,w2/s3,attn --m 640 > gpurun_out/r2_t4_ncu2.log 2>&1
python tools/profile_step.py > gpurun_out/r2_t4_plain3.log 2>&1 &&
ncu --set full --clock-control none --profile-from-start off -k regex:"rmsnorm_mod|cfg_euler_tail|cast_snake|dwconv_ln|conv_out_tanh" -c 12 -o gpurun_out/r2_hbm_kernels -f python tools/profile_step.py > gpurun_out/r2_t4_ncu3.log 2>&1
ls -la gpurun_out/*.ncu-rep
