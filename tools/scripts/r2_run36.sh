# final-kernel evidence of the round: ncu --set full on the GEMMs / attention at both step shapes, launch list of one
# conditioning + CFG step + plain step + decode (each ncu pass only after the same command exited 0 without ncu)
cd /root/repo
python tools/gemm_bench.py --quick --only wo,w2,qkvg,w13,attn --m 1920 > gpurun_out/r2g_plain1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"gemm_pair_kernel|attention_tc_kernel" -c 5 -o gpurun_out/r2g_gemm1920 -f python tools/gemm_bench.py --quick --only wo,w2,qkvg,w13,attn --m 1920 > gpurun_out/r2g_ncu1.log 2>&1
python tools/gemm_bench.py --quick --only wo/s3,w2/s3,qkvg,w13,attn --m 640 > gpurun_out/r2g_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"gemm_pair_kernel|attention_tc_kernel" -c 5 -o gpurun_out/r2g_gemm640 -f python tools/gemm_bench.py --quick --only wo/s3,w2/s3,qkvg,w13,attn --m 640 > gpurun_out/r2g_ncu2.log 2>&1
python tools/profile_step.py > gpurun_out/r2g_plain3.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 2500 --csv --log-file gpurun_out/r2g_launches_step.csv python tools/profile_step.py > gpurun_out/r2g_ncu4.log 2>&1
ls -la gpurun_out | grep r2g
