# evidence re-capture (the first copies were lost with the machine): bench N=1, HBM-kernel ncu, GEMM ncu, cuBLAS probe
cd /root/repo
python bench.py --steps 3 --warmup 3 > gpurun_out/r2_bench_n1_a.json 2> gpurun_out/r2_bench_n1_a.err
python tools/cublas_probe.py > gpurun_out/r2_cublas_plain.log 2>&1 &&
ncu --set full --clock-control none --profile-from-start off -c 8 -o gpurun_out/r2_cublas -f python tools/cublas_probe.py > gpurun_out/r2_cublas_ncu.log 2>&1
python tools/gemm_bench.py --quick --only wo,w2,qkvg,w13,attn --m 1920 > gpurun_out/r2_plain1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"gemm_pair_kernel|attention_tc_kernel" -c 5 -o gpurun_out/r2_gemm1920 -f python tools/gemm_bench.py --quick --only wo,w2,qkvg,w13,attn --m 1920 > gpurun_out/r2_ncu1.log 2>&1
python tools/gemm_bench.py --quick --only wo/s3,w2/s3,qkvg,w13,attn --m 640 > gpurun_out/r2_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"gemm_pair_kernel|attention_tc_kernel" -c 5 -o gpurun_out/r2_gemm640 -f python tools/gemm_bench.py --quick --only wo/s3,w2/s3,qkvg,w13,attn --m 640 > gpurun_out/r2_ncu2.log 2>&1
python tools/profile_step.py > gpurun_out/r2_plain3.log 2>&1 &&
ncu --set full --clock-control none --profile-from-start off -k regex:"rmsnorm_mod|cfg_euler_tail|cast_snake|dwconv_ln|conv_out_tanh" -c 12 -o gpurun_out/r2_hbm_kernels -f python tools/profile_step.py > gpurun_out/r2_ncu3.log 2>&1
python tools/profile_step.py > gpurun_out/r2_plain4.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 2500 --csv --log-file gpurun_out/r2_launches_step.csv python tools/profile_step.py > gpurun_out/r2_ncu4.log 2>&1
ls -la gpurun_out/
