# evidence after the row/TMA epilogue and the conv_out rewrite: launch list of conditioning + CFG step + plain step + decode,
# ncu --set full of the residual-unit GEMMs (k = 1 / k = 7, C = 192) and of conv_out_tanh (each only after the same command
# exited 0 without ncu), then the full GPU suite, the default bench line, the reference arm and smoke()
cd /root/repo
python tools/profile_step.py > gpurun_out/r2h_plain3.log 2>&1 && {
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 2500 --csv --log-file gpurun_out/r2h_launches_step.csv python tools/profile_step.py > gpurun_out/r2h_ncu4.log 2>&1
ncu --set full --clock-control none --profile-from-start off -k regex:conv_out_tanh_kernel -c 1 -o gpurun_out/r2h_hbm_conv_out_tanh_kernel -f python tools/profile_step.py > gpurun_out/r2h_ncu3.log 2>&1
}
python tools/ae_k1_once.py > gpurun_out/r2h_plain5.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"gemm_pair_kernel" -s 2 -c 2 -o gpurun_out/r2h_ae_unit -f python tools/ae_k1_once.py > gpurun_out/r2h_ncu5.log 2>&1
python -m pytest tests -m gpu -x -q -s -p no:cacheprovider 2>&1 | grep -v "^$" | tail -170 > gpurun_out/r2_gpu_tests.log
python bench.py > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/r2_bench_ref.json 2> gpurun_out/r2_bench_ref.err
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2_smoke.log 2>&1
tail -3 gpurun_out/r2_gpu_tests.log; tail -2 gpurun_out/r2_smoke.log; cut -c1-300 gpurun_out/r2_bench_n1.json; cut -c1-300 gpurun_out/r2_bench_ref.json
